// Deterministic synthetic lidar scans for the parity tests and bench: analytic ray casting
// of a spinning multi-beam sensor moving through a procedurally generated street (or a closed room), with the
// sensor pose interpolated per column inside the sweep, Gaussian range noise and optional dropouts / duplicated
// columns.  Test + bench tooling: it produces the identical byte stream that is fed to BOTH the CUDA path and the
// oracle.  No dependency on the product library or on the oracle.
//
// Conventions expected by ImageProjection (reference imageProjection.cpp:199-257): lidar frame x forward, y left,
// z up; points are emitted in firing order = column-major with the head spinning clockwise, so -atan2(y,x) grows
// along the cloud; azimuths sit at image column centres +- 0.3 column of seeded jitter.
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <thread>
#include <vector>

namespace {

struct Box { double lo[3], hi[3]; };
struct Cyl { double cx, cy, r, z0, z1; };

inline uint64_t mix64(uint64_t x) {
    x += 0x9E3779B97F4A7C15ULL; x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ULL; x = (x ^ (x >> 27)) * 0x94D049BB133111EBULL; return x ^ (x >> 31);
}
inline uint64_t hash3(uint64_t a, uint64_t b, uint64_t c) { return mix64(mix64(mix64(a) ^ b) ^ c); }
inline double u01(uint64_t h) { return (double)((h >> 11) + 1) / 9007199254740994.0; }   // (0,1)

struct World {
    std::vector<Box> boxes;
    std::vector<Cyl> cyls;
    // uniform XY grid over the primitives (acceleration only: a ray visits the cells under its ground track in order and
    // stops once its nearest hit lies before the exit of the current cell; the nearest hit is the same minimum the
    // exhaustive loop over all primitives finds)
    double gx0 = 0, gy0 = 0, cell = 4.0;
    int gnx = 0, gny = 0;
    std::vector<int> cellStart, cellItems;   // CSR: items >= 0 box index, < 0 cylinder index -(k+1)
};

void build_room(World& w, uint64_t seed) {
    // closed room x in [-25,25], y in [-15,18], 8 m high walls, 10 poles (SURVEY.md section 8d "closed-room scene")
    const double X0 = -25, X1 = 25, Y0 = -15, Y1 = 18, H = 8, T = 1.0;
    w.boxes.push_back(Box{{X0 - T, Y0 - T, 0}, {X0, Y1 + T, H}});
    w.boxes.push_back(Box{{X1, Y0 - T, 0}, {X1 + T, Y1 + T, H}});
    w.boxes.push_back(Box{{X0, Y0 - T, 0}, {X1, Y0, H}});
    w.boxes.push_back(Box{{X0, Y1, 0}, {X1, Y1 + T, H}});
    for (int i = 0; i < 10; ++i) {
        uint64_t h = hash3(seed, 7001, i);
        double ang = 2 * M_PI * i / 10.0 + 0.3 * u01(h);
        double rad = 6.0 + 7.0 * u01(mix64(h));
        w.cyls.push_back(Cyl{rad * std::cos(ang) * 1.6, rad * std::sin(ang), 0.15 + 0.45 * u01(mix64(h + 1)), 0, 3.0});
    }
    w.boxes.push_back(Box{{8, -9, 0}, {11, -6, 2.0}});
    w.boxes.push_back(Box{{-14, 6, 0}, {-11, 10, 2.5}});
}

void build_street(World& w, uint64_t seed, double x_begin, double x_end) {
    // facades on both sides as rows of buildings with alleys, lamp poles, parked-car blocks, end walls
    for (int side = 0; side < 2; ++side) {
        double sgn = side ? -1.0 : 1.0, x = x_begin; int k = 0;
        while (x < x_end) {
            uint64_t h = hash3(seed, 100 + side, k++);
            double len = 10 + 15 * u01(h), gap = 2 + 3 * u01(mix64(h)), front = 8 + 3 * u01(mix64(h + 1)), hgt = 6 + 9 * u01(mix64(h + 2));
            double y0 = sgn * front, y1 = sgn * (front + 10);
            w.boxes.push_back(Box{{x, std::min(y0, y1), 0}, {x + len, std::max(y0, y1), hgt}});
            x += len + gap;
        }
        x = x_begin + 3; k = 0;
        while (x < x_end) {
            uint64_t h = hash3(seed, 200 + side, k++);
            w.cyls.push_back(Cyl{x, sgn * (6.0 + 0.5 * u01(h)), 0.12 + 0.2 * u01(mix64(h)), 0, 4 + 2 * u01(mix64(h + 1))});
            x += 9 + 6 * u01(mix64(h + 2));
        }
        x = x_begin + 6; k = 0;
        while (x < x_end) {
            uint64_t h = hash3(seed, 300 + side, k++);
            double yc = sgn * (4.6 + 0.4 * u01(h));
            w.boxes.push_back(Box{{x, yc - 0.9, 0}, {x + 4.5, yc + 0.9, 1.5}});
            x += 9 + 14 * u01(mix64(h));
        }
    }
    w.boxes.push_back(Box{{x_begin - 2, -25, 0}, {x_begin - 1, 25, 12}});
    w.boxes.push_back(Box{{x_end + 1, -25, 0}, {x_end + 2, 25, 12}});
}

void build_grid(World& w) {
    const double margin = 0.05;   // primitives are registered a little fat, so a ray grazing a cell corner still meets them
    double x0 = 1e300, y0 = 1e300, x1 = -1e300, y1 = -1e300;
    for (const Box& b : w.boxes) { x0 = std::min(x0, b.lo[0]); y0 = std::min(y0, b.lo[1]); x1 = std::max(x1, b.hi[0]); y1 = std::max(y1, b.hi[1]); }
    for (const Cyl& c : w.cyls) { x0 = std::min(x0, c.cx - c.r); y0 = std::min(y0, c.cy - c.r); x1 = std::max(x1, c.cx + c.r); y1 = std::max(y1, c.cy + c.r); }
    if (x0 > x1) { w.gnx = w.gny = 0; return; }
    w.gx0 = x0 - 1.0; w.gy0 = y0 - 1.0;
    w.gnx = (int)std::ceil((x1 + 1.0 - w.gx0) / w.cell) + 1; w.gny = (int)std::ceil((y1 + 1.0 - w.gy0) / w.cell) + 1;
    std::vector<std::vector<int> > cells((size_t)w.gnx * w.gny);
    auto reg = [&](double ax0, double ay0, double ax1, double ay1, int item) {
        int i0 = (int)std::floor((ax0 - margin - w.gx0) / w.cell), i1 = (int)std::floor((ax1 + margin - w.gx0) / w.cell);
        int j0 = (int)std::floor((ay0 - margin - w.gy0) / w.cell), j1 = (int)std::floor((ay1 + margin - w.gy0) / w.cell);
        i0 = std::max(i0, 0); j0 = std::max(j0, 0); i1 = std::min(i1, w.gnx - 1); j1 = std::min(j1, w.gny - 1);
        for (int j = j0; j <= j1; ++j) for (int i = i0; i <= i1; ++i) cells[(size_t)j * w.gnx + i].push_back(item);
    };
    for (size_t k = 0; k < w.boxes.size(); ++k) reg(w.boxes[k].lo[0], w.boxes[k].lo[1], w.boxes[k].hi[0], w.boxes[k].hi[1], (int)k);
    for (size_t k = 0; k < w.cyls.size(); ++k) reg(w.cyls[k].cx - w.cyls[k].r, w.cyls[k].cy - w.cyls[k].r, w.cyls[k].cx + w.cyls[k].r, w.cyls[k].cy + w.cyls[k].r, -(int)k - 1);
    w.cellStart.assign(cells.size() + 1, 0);
    for (size_t c = 0; c < cells.size(); ++c) w.cellStart[c + 1] = w.cellStart[c] + (int)cells[c].size();
    w.cellItems.resize(w.cellStart.back());
    for (size_t c = 0; c < cells.size(); ++c) std::copy(cells[c].begin(), cells[c].end(), w.cellItems.begin() + w.cellStart[c]);
}

inline double hit_box(const Box& b, const double o[3], const double d[3], double tmax) {
    double t0 = 1e-9, t1 = tmax;
    for (int a = 0; a < 3; ++a) {
        if (std::fabs(d[a]) < 1e-12) { if (o[a] < b.lo[a] || o[a] > b.hi[a]) return -1; continue; }
        double inv = 1.0 / d[a], ta = (b.lo[a] - o[a]) * inv, tb = (b.hi[a] - o[a]) * inv;
        if (ta > tb) std::swap(ta, tb);
        if (ta > t0) t0 = ta;
        if (tb < t1) t1 = tb;
        if (t0 > t1) return -1;
    }
    return t0 > 1e-6 ? t0 : -1;   // origin inside a box: ignored (never happens on the generated paths)
}

inline double hit_cyl(const Cyl& c, const double o[3], const double d[3], double tmax) {
    double ox = o[0] - c.cx, oy = o[1] - c.cy;
    double A = d[0] * d[0] + d[1] * d[1]; if (A < 1e-14) return -1;
    double B = ox * d[0] + oy * d[1], C = ox * ox + oy * oy - c.r * c.r;
    double disc = B * B - A * C; if (disc < 0) return -1;
    double t = (-B - std::sqrt(disc)) / A;
    if (t <= 1e-6 || t >= tmax) return -1;
    double z = o[2] + t * d[2];
    return (z >= c.z0 && z <= c.z1) ? t : -1;
}

// Nearest primitive hit along o + t d with t < best (returns the new best).  Walks the grid cells under the ray in order
// of t (2-D DDA) and tests the primitives registered in each; ends when the nearest hit so far lies before the cell's exit.
inline double nearest_hit(const World& w, const double o[3], const double d[3], double best) {
    if (w.gnx == 0) return best;
    const double cs = w.cell;
    double t0 = 0.0, t1 = best;   // the ground track clipped against the grid rectangle
    const double lo[2] = {w.gx0, w.gy0}, hi[2] = {w.gx0 + cs * w.gnx, w.gy0 + cs * w.gny};
    for (int a = 0; a < 2; ++a) {
        if (std::fabs(d[a]) < 1e-12) { if (o[a] < lo[a] || o[a] >= hi[a]) return best; continue; }
        double inv = 1.0 / d[a], ta = (lo[a] - o[a]) * inv, tb = (hi[a] - o[a]) * inv;
        if (ta > tb) std::swap(ta, tb);
        if (ta > t0) t0 = ta;
        if (tb < t1) t1 = tb;
        if (t0 > t1) return best;
    }
    const double px = o[0] + (t0 + 1e-9) * d[0], py = o[1] + (t0 + 1e-9) * d[1];
    int ix = (int)std::floor((px - w.gx0) / cs), iy = (int)std::floor((py - w.gy0) / cs);
    ix = std::min(std::max(ix, 0), w.gnx - 1); iy = std::min(std::max(iy, 0), w.gny - 1);
    const int sx = d[0] > 0 ? 1 : -1, sy = d[1] > 0 ? 1 : -1;
    const double big = 1e300;
    double tmx = std::fabs(d[0]) < 1e-12 ? big : ((w.gx0 + (ix + (sx > 0 ? 1 : 0)) * cs) - o[0]) / d[0];
    double tmy = std::fabs(d[1]) < 1e-12 ? big : ((w.gy0 + (iy + (sy > 0 ? 1 : 0)) * cs) - o[1]) / d[1];
    const double tdx = std::fabs(d[0]) < 1e-12 ? big : cs / std::fabs(d[0]), tdy = std::fabs(d[1]) < 1e-12 ? big : cs / std::fabs(d[1]);
    for (;;) {
        const size_t c = (size_t)iy * w.gnx + ix;
        for (int k = w.cellStart[c]; k < w.cellStart[c + 1]; ++k) {
            const int it = w.cellItems[k];
            const double t = it >= 0 ? hit_box(w.boxes[it], o, d, best) : hit_cyl(w.cyls[-it - 1], o, d, best);
            if (t > 0 && t < best) best = t;
        }
        const double texit = std::min(tmx, tmy);
        if (best <= texit || texit >= t1) break;
        if (tmx < tmy) { ix += sx; tmx += tdx; if (ix < 0 || ix >= w.gnx) break; }
        else { iy += sy; tmy += tdy; if (iy < 0 || iy >= w.gny) break; }
    }
    return best;
}

}  // namespace

extern "C" {

// All fields plain C; mirrored by a ctypes.Structure in lego_loam_b200/synth.py.
typedef struct {
    int n_scan, horizon_scan;
    double elev_bottom_deg, elev_step_deg;   // beam i elevation = bottom + i*step
    double ang_res_x_deg;                    // image column width (= utility.h ang_res_x for the config)
    double sensor_height;
    int scene;                               // 0 = closed room, 1 = street
    double speed, yaw_amp, yaw_period, sway_amp;   // trajectory
    double noise_sigma, max_range, min_range;
    double dropout_prob, dup_col_prob, jitter_frac;
    double scan_period;
    int traj;                                // 0 = forward drive with sway (above), 1 = closed circle around the origin (loop closure tests)
    double circle_radius;
} lego_synth_cfg;

// Ground-truth sensor pose at time t: out = x,y,z,roll,pitch,yaw (lidar/world frame, z up).
void lego_synth_pose(const lego_synth_cfg* c, double t, double* out6) {
    if (c->traj == 1) {   // counter-clockwise circle, heading along the tangent; one lap takes 2 pi r / speed seconds
        const double r = c->circle_radius, om = c->speed / r, a = om * t;
        out6[0] = r * std::sin(a); out6[1] = -r * std::cos(a); out6[2] = c->sensor_height + 0.02 * std::sin(0.9 * t);
        out6[3] = 0.01 * std::sin(0.7 * t); out6[4] = 0.01 * std::sin(0.5 * t + 1.0); out6[5] = a;
        return;
    }
    double w = 2 * M_PI / c->yaw_period;
    double x = c->speed * t, y = c->sway_amp * std::sin(w * t);
    double yaw = std::atan2(c->sway_amp * w * std::cos(w * t), c->speed > 1e-9 ? c->speed : 1e-9) + c->yaw_amp * std::sin(0.5 * w * t);
    if (c->speed <= 1e-9) yaw = c->yaw_amp * std::sin(0.5 * w * t);
    out6[0] = x; out6[1] = y; out6[2] = c->sensor_height + 0.02 * std::sin(0.9 * t);
    out6[3] = 0.01 * std::sin(0.7 * t); out6[4] = 0.01 * std::sin(0.5 * t + 1.0); out6[5] = yaw;
}

struct lego_synth_world { World w; uint64_t seed; lego_synth_cfg cfg; };

void* lego_synth_create(const lego_synth_cfg* c, uint64_t seed, double total_time) {
    lego_synth_world* s = new lego_synth_world();
    s->seed = seed; s->cfg = *c;
    if (c->scene == 0) build_room(s->w, seed);
    else build_street(s->w, seed, -60.0, c->speed * total_time + 120.0);
    build_grid(s->w);
    return s;
}
void lego_synth_destroy(void* h) { delete static_cast<lego_synth_world*>(h); }

// One scan. xyzi: cap*fpp floats (fpp = 4: x y z intensity, 3: x y z packed), ring: cap u16. Returns the number of points
// written (<= cap), firing order.
int lego_synth_scan_fmt(void* h, int scan_idx, float* xyzi, uint16_t* ring, int cap, int fpp) {
    const lego_synth_world& S = *static_cast<lego_synth_world*>(h);
    const lego_synth_cfg& c = S.cfg;
    const int R = c.n_scan, C = c.horizon_scan;
    const double t_scan = scan_idx * c.scan_period;
    int n = 0;
    std::vector<double> ce(R), se(R);
    for (int i = 0; i < R; ++i) { double e = (c.elev_bottom_deg + i * c.elev_step_deg) * M_PI / 180.0; ce[i] = std::cos(e); se[i] = std::sin(e); }
    for (int k = 0; k < C; ++k) {
        const int col = C - 1 - k;   // clockwise: azimuth decreases along the firing order
        uint64_t hc = hash3(S.seed, 0x5000 + scan_idx, k);
        int reps = (c.dup_col_prob > 0 && u01(mix64(hc + 3)) < c.dup_col_prob) ? 2 : 1;
        for (int rep = 0; rep < reps; ++rep) {
            double jit = (2.0 * u01(mix64(hc + 11 * rep)) - 1.0) * c.jitter_frac;
            double az = ((col - C / 2) + jit) * c.ang_res_x_deg * M_PI / 180.0;
            double frac = (k + 0.5 * rep) / (double)C;
            double P[6]; lego_synth_pose(&c, t_scan + frac * c.scan_period, P);
            double cr = std::cos(P[3]), sr = std::sin(P[3]), cp = std::cos(P[4]), sp = std::sin(P[4]), cy = std::cos(P[5]), sy = std::sin(P[5]);
            // world_from_sensor = Rz(yaw) * Ry(pitch) * Rx(roll)
            double Rm[3][3] = {{cy * cp, cy * sp * sr - sy * cr, cy * sp * cr + sy * sr},
                               {sy * cp, sy * sp * sr + cy * cr, sy * sp * cr - cy * sr},
                               {-sp, cp * sr, cp * cr}};
            double ca = std::cos(az), sa = std::sin(az);
            for (int i = 0; i < R; ++i) {
                double ds[3] = {ce[i] * ca, ce[i] * sa, se[i]};
                double d[3], o[3] = {P[0], P[1], P[2]};
                for (int a = 0; a < 3; ++a) d[a] = Rm[a][0] * ds[0] + Rm[a][1] * ds[1] + Rm[a][2] * ds[2];
                double best = c.max_range;
                if (d[2] < -1e-9) { double t = -o[2] / d[2]; if (t < best) best = t; }
                best = nearest_hit(S.w, o, d, best);
                if (best >= c.max_range) continue;   // no return
                uint64_t hr = hash3(S.seed ^ 0xABCDEF, ((uint64_t)scan_idx << 24) ^ (uint64_t)(k * 2 + rep), i);
                if (c.dropout_prob > 0 && u01(mix64(hr + 5)) < c.dropout_prob) continue;
                double u1 = u01(hr), u2 = u01(mix64(hr + 1));
                double g = std::sqrt(-2.0 * std::log(u1)) * std::cos(2 * M_PI * u2);
                double rng = best + c.noise_sigma * g;
                if (rng < c.min_range) continue;
                if (n >= cap) return n;
                float* q = xyzi + (size_t)fpp * n;
                q[0] = (float)(rng * ds[0]); q[1] = (float)(rng * ds[1]); q[2] = (float)(rng * ds[2]); if (fpp == 4) q[3] = 10.0f;
                ring[n] = (uint16_t)i; ++n;
            }
        }
    }
    return n;
}

int lego_synth_scan(void* h, int scan_idx, float* xyzi, uint16_t* ring, int cap) { return lego_synth_scan_fmt(h, scan_idx, xyzi, ring, cap, 4); }

// count scans starting at scan0, multi-threaded, written straight into a strided destination: scan s goes to
// xyz + s*scan_stride_points*fpp floats / ring + s*scan_stride_points (lets a caller interleave several sequences in one
// pinned [scan][sequence][cap] buffer without an intermediate copy).  npts advances by npts_stride per scan.
void lego_synth_batch_fmt(void* h, int scan0, int count, float* xyz, uint16_t* ring, int* npts, int cap, int threads, int fpp,
                          long scan_stride_points, int npts_stride) {
    if (threads < 1) threads = 1;
    std::vector<std::thread> pool;
    for (int t = 0; t < threads; ++t)
        pool.emplace_back([=]() {
            for (int s = t; s < count; s += threads)
                npts[(size_t)s * npts_stride] = lego_synth_scan_fmt(h, scan0 + s, xyz + (size_t)s * scan_stride_points * fpp, ring + (size_t)s * scan_stride_points, cap, fpp);
        });
    for (auto& th : pool) th.join();
}

// The same with independent strides for the two channels (in floats and in uint16 elements): lets a caller keep a scan's ring
// channel in the same slot as its points.
void lego_synth_batch_strided(void* h, int scan0, int count, float* xyz, uint16_t* ring, int* npts, int cap, int threads, int fpp,
                              long xyz_stride_floats, long ring_stride_elems, int npts_stride) {
    if (threads < 1) threads = 1;
    std::vector<std::thread> pool;
    for (int t = 0; t < threads; ++t)
        pool.emplace_back([=]() {
            for (int s = t; s < count; s += threads)
                npts[(size_t)s * npts_stride] = lego_synth_scan_fmt(h, scan0 + s, xyz + (size_t)s * xyz_stride_floats, ring + (size_t)s * ring_stride_elems, cap, fpp);
        });
    for (auto& th : pool) th.join();
}

// count scans starting at scan0, multi-threaded. xyzi: count*cap*4, ring: count*cap, npts: count.
void lego_synth_batch(void* h, int scan0, int count, float* xyzi, uint16_t* ring, int* npts, int cap, int threads) {
    if (threads < 1) threads = 1;
    std::vector<std::thread> pool;
    for (int t = 0; t < threads; ++t)
        pool.emplace_back([=]() {
            for (int s = t; s < count; s += threads)
                npts[s] = lego_synth_scan(h, scan0 + s, xyzi + (size_t)s * cap * 4, ring + (size_t)s * cap, cap);
        });
    for (auto& th : pool) th.join();
}

}  // extern "C"
