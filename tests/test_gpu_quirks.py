"""GPU micro-tests of the reference's quirks (SURVEY.md section 8, quirk list Q1-Q17): every test CONSTRUCTS an input that
forces one quirk, asserts on the reference's own output (oracle/_ref) that the quirk really fired (count > 0) and then
compares the CUDA path with it bit for bit.  Random scenes only trigger these cases by accident; here they are certain.

The constructed world: a flat ground plane and a cylindrical wall of radius 30 m around a 16-ring sensor 0.7 m above the
ground (every beam returns: rings 0..6 hit the ground, rings 7..15 the wall), one return per image cell at the cell's
centre, in firing order (column-major, clockwise).  Tests then add, move or drop individual returns."""
import numpy as np
import pytest

import lego_loam_b200  # noqa: F401
from conftest import require_ref
from lego_loam_b200.capi import Pipeline
from lego_loam_b200.params import make_params

pytestmark = pytest.mark.gpu

R, C = 16, 1800
ELEV = np.deg2rad(-15.0 + 2.0 * np.arange(R))
H, WALL = 0.7, 30.0


def base_ranges(noise=None):
    """range[ring, col] of the flat world"""
    rng = np.empty((R, C))
    for i in range(R):
        wall = WALL / np.cos(ELEV[i])
        ground = H / np.sin(-ELEV[i]) if ELEV[i] < 0 else np.inf
        rng[i, :] = min(wall, ground)
    if noise is not None:
        rng = rng + noise
    return rng


def cloud_from_ranges(rng, keep=None, col_shift=0.0):
    """-> xyzi float32[P, 4], ring uint16[P] in firing order; keep[ring, col] = False drops a return"""
    pts, rings = [], []
    for k in range(C):
        col = C - 1 - k                                  # clockwise: -atan2(y, x) grows along the cloud
        az = np.deg2rad(((col - C // 2) + col_shift) * 0.2)
        for i in range(R):
            if keep is not None and not keep[i, col]:
                continue
            r = rng[i, col]
            pts.append((r * np.cos(ELEV[i]) * np.cos(az), r * np.cos(ELEV[i]) * np.sin(az), r * np.sin(ELEV[i]), 10.0))
            rings.append(i)
    return np.asarray(pts, np.float32), np.asarray(rings, np.uint16)


def both(sensor="vlp16"):
    import refpy
    require_ref(sensor)
    return Pipeline(make_params(sensor), lanes=1, max_points=R * C + 4096), refpy.RefPipeline(sensor)


def feed(pl, ref, xyzi, ring, k=0, stages=3):
    rf = ref.feed(xyzi, ring, 0.1 * k, stages=stages)
    cap = R * C + 4096
    x = np.zeros((1, cap, 4), np.float32); x[0, :len(xyzi)] = xyzi
    r = None
    if ring is not None:
        r = np.zeros((1, cap), np.uint16); r[0, :len(ring)] = ring
    pl.step(x, r, [len(xyzi)], [0.1 * k], stage_mask=stages)
    return rf


def bits(a):
    return np.ascontiguousarray(a, np.float32).view(np.uint32)


def same_images(pl, ref):
    """every ImageProjection output bit for bit (the comparison of tests/test_gpu_image_projection.py)"""
    from test_gpu_image_projection import _compare_ip
    errs = _compare_ip(pl, ref)
    assert not errs, errs


# ---------------------------------------------------------------------------------------------------- ImageProjection
def test_q3_cell_collisions_last_point_in_input_order_wins():
    """imageProjection.cpp:248-255: a later point overwrites rangeMat / fullCloud of its cell.  400 cells get a second return:
    200 of them LATER in the cloud (it must win), 200 EARLIER (it must lose)."""
    rng = base_ranges()
    xyzi, ring = cloud_from_ranges(rng)
    cols = np.arange(100, 1700, 4)
    extra = []
    for n, col in enumerate(cols):
        az = np.deg2rad((col - C // 2) * 0.2)
        r = 12.0 + 0.01 * n
        extra.append((r * np.cos(ELEV[10]) * np.cos(az), r * np.cos(ELEV[10]) * np.sin(az), r * np.sin(ELEV[10]), 10.0))
    extra = np.asarray(extra, np.float32)
    late, early = extra[:200], extra[200:]
    cloud = np.concatenate([xyzi[:1], early, xyzi[1:-1], late, xyzi[-1:]])      # first / last point (orientation) untouched
    rings = np.concatenate([ring[:1], np.full(200, 10, np.uint16), ring[1:-1], np.full(200, 10, np.uint16), ring[-1:]])
    pl, ref = both()
    feed(pl, ref, cloud, rings, stages=1)
    rm = ref.get("rangeMat").reshape(R, C)
    r_extra = np.sqrt((extra[:, :3].astype(np.float64) ** 2).sum(1))
    won = np.abs(rm[10, cols] - r_extra) < 1e-3
    assert won[:200].sum() == 200 and won[200:].sum() == 0, "the quirk did not fire as constructed: %d / %d" % (won[:200].sum(), won[200:].sum())
    same_images(pl, ref)
    pl.close(); ref.close()


def test_q4_ground_minus_one_overwrites_a_one():
    """imageProjection.cpp:273-278 vs 286-289: in a column whose ring-2 return is missing, the pair (0, 1) marks both cells
    ground (1), then the pair (1, 2) finds its upper cell invalid and writes -1 over ring 1's 1: that cell is NOT ground,
    gets label 0 and takes part in the segmentation."""
    rng = base_ranges()
    keep = np.ones((R, C), bool)
    cols = np.arange(50, 1750, 7)
    keep[2, cols] = False
    xyzi, ring = cloud_from_ranges(rng, keep)
    pl, ref = both()
    feed(pl, ref, xyzi, ring, stages=1)
    gm = ref.get("groundMat").reshape(R, C)
    fired = int(((gm[1, cols] == -1) & (gm[0, cols] == 1)).sum())
    assert fired == len(cols), fired
    assert (gm[1, np.setdiff1d(np.arange(C), cols)] == 1).all()
    same_images(pl, ref)
    pl.close(); ref.close()


def _blob(rng, rows_cols, r=10.0):
    for i, j in rows_cols:
        rng[i, j] = r / np.cos(ELEV[i])


def test_q5_q6_seed_row_not_counted_and_label_numbering():
    """imageProjection.cpp:376-384 vs 431: the BFS seed's own row only counts towards segmentValidLineNum when ANOTHER cell
    of that row is pushed.  Blob A: 1 + 2 + 2 cells on three rows, seed alone in its row -> 2 counted rows -> invalid (999999)
    although it spans 3 rows.  Blob B: 2 + 2 + 2 cells -> 3 counted rows -> valid.  Valid segments are numbered 1, 2, ... in
    row-major order of their seeds (Q6)."""
    rng = base_ranges()
    blobs_a = [[(9, c), (10, c), (10, c + 1), (11, c), (11, c + 1)] for c in range(200, 1600, 100)]
    blobs_b = [[(9, c), (9, c + 1), (10, c), (10, c + 1), (11, c), (11, c + 1)] for c in range(250, 1600, 100)]
    for b in blobs_a + blobs_b:
        _blob(rng, b)
    xyzi, ring = cloud_from_ranges(rng)
    pl, ref = both()
    feed(pl, ref, xyzi, ring, stages=1)
    lm = ref.get("labelMat").reshape(R, C)
    a_invalid = sum(all(lm[i, j] == 999999 for i, j in b) for b in blobs_a)
    b_labels = [lm[b[0]] for b in blobs_b]
    assert a_invalid == len(blobs_a), "seed-row quirk did not fire: %d of %d" % (a_invalid, len(blobs_a))
    assert all(0 < l < 999999 for l in b_labels) and all(all(lm[i, j] == lm[b[0]] for i, j in b) for b in blobs_b)
    assert list(b_labels) == sorted(b_labels) and len(set(b_labels)) == len(b_labels)       # numbered in seed order
    same_images(pl, ref)
    pl.close(); ref.close()


def test_q2_row_truncation_and_column_wrap_without_ring_channel():
    """imageProjection.cpp:214, 229-242: rowIdn / columnIdn are size_t.  Without the ring channel a return 0.3 deg BELOW the
    lowest beam gives (angle + ang_bottom) / ang_res_y = -0.1, which truncates to row 0 (the `< 0` test is dead code); the
    column needs exactly one wrap for a quarter of the image."""
    rng = base_ranges()
    keep = np.ones((R, C), bool)
    keep[0, :] = False                                   # remove ring 0, then re-add it 0.3 deg lower
    xyzi, ring = cloud_from_ranges(rng, keep)
    low = np.deg2rad(-15.3)
    pts = []
    for k in range(C):
        col = C - 1 - k
        az = np.deg2rad((col - C // 2) * 0.2)
        r = H / np.sin(-low)
        pts.append((k, (r * np.cos(low) * np.cos(az), r * np.cos(low) * np.sin(az), r * np.sin(low), 10.0)))
    # interleave in firing order: the low return leads its column
    out = []
    per_col = R - 1
    for k in range(C):
        out.append(pts[k][1])
        out.extend(map(tuple, xyzi[k * per_col:(k + 1) * per_col]))
    cloud = np.asarray(out, np.float32)
    pl, ref = both("vlp16_noring")
    feed(pl, ref, cloud, None, stages=1)
    rm = ref.get("rangeMat").reshape(R, C)
    filled = int((rm[0] < 1e30).sum())
    assert filled == C, "row 0 was not filled by the truncated rows: %d" % filled
    same_images(pl, ref)
    pl.close(); ref.close()


# ------------------------------------------------------------------------------------------------ FeatureAssociation
def same_features(pl, ref, tag=""):
    adj_g, adj_r = pl.download("segmentedCloudAdj"), ref.cloud("segmentedCloudAdj")
    assert adj_g.shape == adj_r.shape and np.array_equal(bits(adj_g), bits(adj_r)), tag + "segmentedCloudAdj"
    S = len(adj_r)
    assert np.array_equal(bits(pl.download("cloudCurvature")[5:S - 5]), bits(ref.get("cloudCurvature")[5:S - 5])), tag + "cloudCurvature"
    assert np.array_equal(pl.download("cloudNeighborPicked"), ref.get("cloudNeighborPicked")), tag + "cloudNeighborPicked"
    assert np.array_equal(pl.download("cloudLabel"), ref.get("cloudLabel")), tag + "cloudLabel"
    for name in ("cornerPointsSharp", "cornerPointsLessSharp", "surfPointsFlat"):
        g, w = pl.download(name), ref.cloud(name)
        assert g.shape == w.shape and np.array_equal(bits(g), bits(w)), tag + name
    g, w = pl.download("surfPointsLessFlat"), ref.cloud("surfPointsLessFlat")
    assert g.shape == w.shape and np.allclose(g, w, rtol=1e-5, atol=1e-6), tag + "surfPointsLessFlat"
    for name in ("transformCur", "transformSum"):
        assert np.max(np.abs(pl.download(name) - ref.get(name))) < 1e-4, tag + name


def _sectors(ref):
    """(sp, ep) of every non-empty sector, featureAssociation.cpp:690-697"""
    st, en = ref.get("startRingIndex"), ref.get("endRingIndex")
    out = []
    for i in range(R):
        for j in range(6):
            sp = (int(st[i]) * (6 - j) + int(en[i]) * j) // 6
            ep = (int(st[i]) * (5 - j) + int(en[i]) * (j + 1)) // 6 - 1
            if sp < ep:
                out.append((sp, ep))
    return out


def _bumpy_world(seed, quantum=None):
    """the flat world with boxes of different depth in front of the wall (edges for the corner features) and ripples on
    the wall; quantum rounds every range to a grid, which makes equal curvatures (std::sort ties) common"""
    rs = np.random.default_rng(seed)
    rng = base_ranges()
    for c0 in range(40, 1760, 60):
        w = int(rs.integers(6, 30)); depth = float(rs.uniform(8.0, 25.0))
        for i in range(7, R):
            rng[i, c0:c0 + w] = depth / np.cos(ELEV[i])
    rng[7:] += rs.normal(0.0, 0.02, (R - 7, C))
    rng[:7] += rs.normal(0.0, 0.005, (7, C))
    if quantum:
        rng = np.round(rng / quantum) * quantum
    return rng


def test_q10_q11_q12_sector_scan_bounds_and_stale_slot():
    """featureAssociation.cpp:699-702, 713-718, 746-749 and 210-223, 624: std::sort covers [sp, ep) but both selection passes
    also visit slot ep (Q10); the flat pass stops right after its 4th pick, before suppressing its neighbours (Q12); slot 4
    of cloudSmoothness and the first five work-array entries are never written, so with zeroed memory scan 0 picks segmented
    point 0 as its first flat feature through slot 4 = {0, index 0} (Q11)."""
    pl, ref = both()
    fired_ep = fired_4 = 0
    for k in range(3):
        xyzi, ring = cloud_from_ranges(_bumpy_world(100 + k))
        feed(pl, ref, xyzi, ring, k, stages=3)
        lab = ref.get("cloudLabel")
        if k == 0:
            assert lab[0] == -1, "Q11 did not fire: segmented point 0 is not a flat feature of scan 0 (label %d)" % lab[0]
        for sp, ep in _sectors(ref):
            fired_ep += int(lab[ep] != 0)
            fired_4 += int((lab[sp:ep + 1] == -1).sum() == 4)
        same_features(pl, ref, "scan %d: " % k)
    assert fired_ep > 0 and fired_4 > 0, "Q10 fired %d times, Q12 %d times" % (fired_ep, fired_4)
    pl.close(); ref.close()


def test_q17_std_sort_ties_decide_features():
    """utility.h:144-148 + featureAssociation.cpp:699: by_value compares only the curvature and std::sort is unstable, so
    equal curvatures are ordered by libstdc++'s introsort.  Ranges on a 1 cm grid make most sectors hold ties."""
    pl, ref = both()
    tied_sectors = 0
    for k in range(4):
        xyzi, ring = cloud_from_ranges(_bumpy_world(200 + k, quantum=0.01))
        feed(pl, ref, xyzi, ring, k, stages=3)
        curv = ref.get("cloudCurvature")
        for sp, ep in _sectors(ref):
            c = curv[sp:ep]
            tied_sectors += int(len(np.unique(c)) < len(c))
        same_features(pl, ref, "scan %d: " % k)
    assert tied_sectors > 100, "only %d sectors with equal curvatures" % tied_sectors
    pl.close(); ref.close()


def test_q9_one_half_passed_flag_for_the_whole_cloud():
    """featureAssociation.cpp:493-520: halfPassed is set once, by the first ring, and never reset, so every later ring runs the
    second branch from its first point on.  A cloud whose first return sits 100 columns into the sweep (start orientation
    shifted) moves the place where the flag flips; relative times must still agree bit for bit."""
    pl, ref = both()
    for k, shift in enumerate((0, 100, 900)):
        rng = _bumpy_world(300 + k)
        xyzi, ring = cloud_from_ranges(rng)
        cut = shift * R
        xyzi, ring = np.concatenate([xyzi[cut:], xyzi[:cut]]), np.concatenate([ring[cut:], ring[:cut]])
        feed(pl, ref, xyzi, ring, k, stages=3)
        adj = ref.cloud("segmentedCloudAdj")
        rel = (adj[:, 3] - np.floor(adj[:, 3])) / 0.1
        assert rel.min() >= -0.05 and rel.max() > 0.9, "relative times do not span the sweep: %g .. %g" % (rel.min(), rel.max())
        same_features(pl, ref, "shift %d: " % shift)
    pl.close(); ref.close()


# ------------------------------------------------------------------------------------------------------------ mapping
def test_q16_mapping_cadence_every_fourth_scan():
    """featureAssociation.cpp:284, 314, 1790-1794 (clouds every 2nd scan) and mapOptmization.cpp:1499 (>= 0.3 s): at 10 Hz the
    mapping body runs on every 4th scan (the first one as soon as corner, surf and outlier clouds of one stamp exist)."""
    pl, ref = both()
    ran = []
    for k in range(24):
        xyzi, ring = cloud_from_ranges(_bumpy_world(400))        # a static world: the poses stay at zero
        rf = feed(pl, ref, xyzi, ring, k, stages=7)
        poses, flags = pl.poses()
        assert bool(flags[0] & 8) == bool(rf & 8), "scan %d" % k
        if rf & 8:
            ran.append(k)
        assert np.max(np.abs(poses[0, 6:12] - ref.get("transformAftMapped"))) < 1e-4
    assert len(ran) == 6 and ran[0] <= 2 and all(b - a == 4 for a, b in zip(ran, ran[1:])), ran
    pl.close(); ref.close()


def test_q14_degenerate_direction_is_projected_out():
    """featureAssociation.cpp:1329-1356, 1677-1680 and mapOptmization.cpp:1278-1305: in a world that does not change along
    the direction of travel (flat ground, a wall at constant distance all around, returns at the same ranges while the
    sensor "moves") the translation along the ground plane is unobservable: eigenvalues below the threshold zero rows of
    matP, which is kept across calls.  isDegenerate must be set on both sides and the poses must still agree."""
    pl, ref = both()
    degenerate_calls = 0
    for k in range(14):
        rs = np.random.default_rng(500 + k)
        rng = base_ranges()
        rng[7:] += rs.normal(0.0, 0.01, (R - 7, C)); rng[:7] += rs.normal(0.0, 0.003, (7, C))
        xyzi, ring = cloud_from_ranges(rng)
        rf = feed(pl, ref, xyzi, ring, k, stages=7)
        poses, flags = pl.poses()
        if rf & 8:
            d_ref, d_dev = int(ref.get("moDegenerate")[0]), int(pl.download("moState")[0])
            assert d_ref == d_dev, "scan %d: isDegenerate %d vs %d" % (k, d_ref, d_dev)
            degenerate_calls += d_ref
        assert np.max(np.abs(poses[0, :6] - ref.get("transformSum"))) < 1e-4, "scan %d transformSum" % k
        assert np.max(np.abs(poses[0, 6:12] - ref.get("transformAftMapped"))) < 1e-4, "scan %d transformAftMapped" % k
    assert degenerate_calls > 0, "the scene did not make scan-to-map degenerate"
    pl.close(); ref.close()
