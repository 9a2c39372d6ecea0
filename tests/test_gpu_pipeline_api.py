"""GPU tests of the batched pipeline's C ABI beyond plain parity: the per-lane pointer form of the step
(lego_pipeline_step_lanes, x y z packed or x y z intensity, host or device memory), the input-lifetime call
(lego_pipeline_input_wait) and the capacity errors the header promises (include/lego_b200.h)."""
import numpy as np
import pytest

import lego_loam_b200  # noqa: F401
from lego_loam_b200.capi import LegoError, Pipeline, load
from lego_loam_b200.params import make_params
from lego_loam_b200.synth import SynthSequence

pytestmark = pytest.mark.gpu

NAMES = ("labelMat", "segmentedCloud", "cornerPointsSharp", "cornerPointsLessSharp", "surfPointsFlat", "surfPointsLessFlat")


def _scans(sensor, seeds, n, speed=4.0):
    seqs = [SynthSequence(sensor, seed=s, scene="street", speed=speed) for s in seeds]
    return seqs, [[q.scan(k) for q in seqs] for k in range(n)]


def _same(a, b):
    return a.shape == b.shape and np.array_equal(a.view(np.uint8), b.view(np.uint8))


def test_step_lanes_xyz_and_xyzi_equal_padded_step():
    """Three forms of the same 10 scans (2 lanes) give bit-identical images, features and poses: the padded xyzi block of
    lego_pipeline_step, one xyzi pointer per lane, one packed xyz pointer per lane (the input intensity is never read,
    imageProjection.cpp:248-255)."""
    sensor = "vlp16"
    seqs, scans = _scans(sensor, (31, 32), 10)
    cap = seqs[0].cap
    pls = [Pipeline(make_params(sensor), lanes=2, max_points=cap) for _ in range(3)]
    for k, pair in enumerate(scans):
        x = np.zeros((2, cap, 4), np.float32); r = np.zeros((2, cap), np.uint16)
        n = np.array([len(p[0]) for p in pair], np.int32)
        for l, (xyzi, ring) in enumerate(pair):
            x[l, :n[l]] = xyzi; r[l, :n[l]] = ring
        st = [k * 0.1, k * 0.1]
        pls[0].step(x, r, n, st)
        xi = [np.ascontiguousarray(p[0]) for p in pair]
        x3 = [np.ascontiguousarray(p[0][:, :3]) for p in pair]
        rg = [np.ascontiguousarray(p[1]) for p in pair]
        rp = np.array([a.ctypes.data for a in rg], np.uint64)
        pls[1].step_lanes(np.array([a.ctypes.data for a in xi], np.uint64), rp, n, st, floats_per_point=4)
        pls[2].step_lanes(np.array([a.ctypes.data for a in x3], np.uint64), rp, n, st, floats_per_point=3)
        for pl in pls[1:]:
            pl.input_wait()          # the numpy buffers go out of scope at the end of this iteration
        p0 = pls[0].poses()[0]
        for pl in pls[1:]:
            assert np.array_equal(pl.poses()[0], p0), "poses differ at scan %d" % k
            for lane in range(2):
                for name in NAMES:
                    assert _same(pl.download(name, lane), pls[0].download(name, lane)), "%s lane %d scan %d" % (name, lane, k)
    assert np.abs(p0[:, 6:12]).max() > 0.05
    for pl in pls:
        pl.close()


def test_step_lanes_from_device_memory():
    import torch
    sensor = "vlp16"
    seqs, scans = _scans(sensor, (41,), 6)
    cap = seqs[0].cap
    a = Pipeline(make_params(sensor), lanes=1, max_points=cap)
    b = Pipeline(make_params(sensor), lanes=1, max_points=cap)
    for k, pair in enumerate(scans):
        xyzi, ring = pair[0]
        n = np.array([len(xyzi)], np.int32)
        x3 = np.ascontiguousarray(xyzi[:, :3])
        a.step_lanes(np.array([x3.ctypes.data], np.uint64), np.array([ring.ctypes.data], np.uint64), n, [k * 0.1], floats_per_point=3)
        a.input_wait()
        dx = torch.from_numpy(x3).cuda(); dr = torch.from_numpy(ring.view(np.int16)).cuda()
        torch.cuda.synchronize()
        b.step_lanes(np.array([dx.data_ptr()], np.uint64), np.array([dr.data_ptr()], np.uint64), n, [k * 0.1], floats_per_point=3, on_device=True)
        b.input_wait()               # dx / dr may be freed once the kernels that read them have run
        assert np.array_equal(a.poses()[0], b.poses()[0])
        assert _same(a.download("labelMat"), b.download("labelMat"))
    a.close(); b.close()


def test_step_lanes_one_block_per_lane():
    """A lane whose ring channel starts exactly where its points end is uploaded with one copy (include/lego_b200.h); the
    results equal the two-copy path's.  Mixed in one step: lane 0 adjacent, lane 1 separate buffers."""
    sensor = "vlp16"
    seqs, scans = _scans(sensor, (61, 62), 7)
    cap = seqs[0].cap
    a = Pipeline(make_params(sensor), lanes=2, max_points=cap)
    b = Pipeline(make_params(sensor), lanes=2, max_points=cap)
    for k, pair in enumerate(scans):
        n = np.array([len(p[0]) for p in pair], np.int32)
        x3 = [np.ascontiguousarray(p[0][:, :3]) for p in pair]
        rg = [np.ascontiguousarray(p[1]) for p in pair]
        blob = np.zeros(n[0] * 14, np.uint8)
        blob[:n[0] * 12] = x3[0].view(np.uint8).reshape(-1); blob[n[0] * 12:] = rg[0].view(np.uint8)
        st = [k * 0.1] * 2
        a.step_lanes(np.array([x.ctypes.data for x in x3], np.uint64), np.array([r.ctypes.data for r in rg], np.uint64), n, st, floats_per_point=3)
        b.step_lanes(np.array([blob.ctypes.data, x3[1].ctypes.data], np.uint64), np.array([blob.ctypes.data + int(n[0]) * 12, rg[1].ctypes.data], np.uint64),
                     n, st, floats_per_point=3)
        a.input_wait(); b.input_wait()
        assert np.array_equal(a.poses()[0], b.poses()[0]), "scan %d" % k
        for lane in range(2):
            for name in NAMES:
                assert _same(a.download(name, lane), b.download(name, lane)), "%s lane %d scan %d" % (name, lane, k)
    a.close(); b.close()


def test_step_lanes_runs_of_lanes_share_a_copy():
    """Lanes whose blocks follow one another in host memory travel in one copy (include/lego_b200.h): with src_on_device = 3 the
    slots of one arena (fixed slot size, unused tails between the lanes), with src_on_device = 0 only blocks that touch
    exactly -- here with even and odd point counts, so that some joints are 4-byte aligned and some are not.  Same results as
    separate buffers, x y z and x y z intensity."""
    sensor = "vlp16"
    seqs, scans = _scans(sensor, (71, 72, 73, 74), 6)
    cap = seqs[0].cap
    for fpp in (3, 4):
        bpp = 4 * fpp + 2
        a = Pipeline(make_params(sensor), lanes=4, max_points=cap)
        b = Pipeline(make_params(sensor), lanes=4, max_points=cap)
        c = Pipeline(make_params(sensor), lanes=4, max_points=cap)
        slot = cap * bpp
        arena = np.zeros(4 * slot + 64, np.uint8)
        a0 = (-arena.ctypes.data) % 16                   # slot 0 starts 16-byte aligned
        tight = np.zeros(4 * slot + 64, np.uint8)
        t0 = (-tight.ctypes.data) % 16
        for k, four in enumerate(scans):
            n = np.array([len(p[0]) - (l + k) % 2 for l, p in enumerate(four)], np.int32)
            xs = [np.ascontiguousarray(p[0][:n[l], :fpp]) for l, p in enumerate(four)]
            rs = [np.ascontiguousarray(p[1][:n[l]]) for l, p in enumerate(four)]
            xa, ra, xt, rt = [], [], [], []
            cur = t0
            for l in range(4):
                raw = np.concatenate([xs[l].view(np.uint8).reshape(-1), rs[l].view(np.uint8).reshape(-1)])
                o = a0 + l * slot
                arena[o:o + len(raw)] = raw
                arena[o + len(raw):o + slot] = 0xAB                      # the tail of a slot is never interpreted
                xa.append(arena.ctypes.data + o); ra.append(arena.ctypes.data + o + int(n[l]) * 4 * fpp)
                tight[cur:cur + len(raw)] = raw
                xt.append(tight.ctypes.data + cur); rt.append(tight.ctypes.data + cur + int(n[l]) * 4 * fpp)
                cur += len(raw)                                          # the next lane touches this one
            st = [k * 0.1] * 4
            a.step_lanes(np.array([x.ctypes.data for x in xs], np.uint64), np.array([r.ctypes.data for r in rs], np.uint64), n, st, floats_per_point=fpp)
            b.step_lanes(np.array(xa, np.uint64), np.array(ra, np.uint64), n, st, floats_per_point=fpp, on_device=3)
            c.step_lanes(np.array(xt, np.uint64), np.array(rt, np.uint64), n, st, floats_per_point=fpp, on_device=0)
            a.input_wait(); b.input_wait(); c.input_wait()
            for other, what in ((b, "arena"), (c, "touching")):
                assert np.array_equal(a.poses()[0], other.poses()[0]), "%s fpp %d scan %d" % (what, fpp, k)
                for lane in range(4):
                    for name in NAMES:
                        assert _same(a.download(name, lane), other.download(name, lane)), "%s %s lane %d scan %d" % (what, name, lane, k)
        a.close(); b.close(); c.close()


def test_step_lanes_from_page_locked_memory():
    """src_on_device = 2: the lanes' scans sit in page-locked host memory and one packing kernel gathers them (no per-lane
    copies).  Same results as ordinary host memory, with odd point counts and unaligned lane pointers."""
    import torch
    sensor = "vlp16"
    seqs, scans = _scans(sensor, (51, 52, 53), 7)
    cap = seqs[0].cap
    a = Pipeline(make_params(sensor), lanes=3, max_points=cap)
    b = Pipeline(make_params(sensor), lanes=3, max_points=cap)
    pin_x = torch.zeros((3, cap * 3 + 8), dtype=torch.float32).pin_memory()
    pin_r = torch.zeros((3, cap + 8), dtype=torch.int16).pin_memory()
    for k, trio in enumerate(scans):
        n = np.array([len(p[0]) - (l + k) % 2 for l, p in enumerate(trio)], np.int32)       # odd and even counts
        xs, rs, xp, rp = [], [], [], []
        for l, (xyzi, ring) in enumerate(trio):
            x3 = np.ascontiguousarray(xyzi[:n[l], :3]); rg = np.ascontiguousarray(ring[:n[l]])
            xs.append(x3); rs.append(rg)
            off = l                                         # lane l starts l floats / l uint16 into its row: 4- and 2-byte aligned only
            pin_x[l, off:off + 3 * n[l]] = torch.from_numpy(x3.reshape(-1))
            pin_r[l, off:off + n[l]] = torch.from_numpy(rg.view(np.int16))
            xp.append(pin_x[l].data_ptr() + 4 * off); rp.append(pin_r[l].data_ptr() + 2 * off)
        st = [k * 0.1] * 3
        a.step_lanes(np.array([x.ctypes.data for x in xs], np.uint64), np.array([r.ctypes.data for r in rs], np.uint64), n, st, floats_per_point=3)
        b.step_lanes(np.array(xp, np.uint64), np.array(rp, np.uint64), n, st, floats_per_point=3, on_device=2)
        a.input_wait(); b.input_wait()
        assert np.array_equal(a.poses()[0], b.poses()[0]), "scan %d" % k
        for lane in range(3):
            for name in NAMES:
                assert _same(a.download(name, lane), b.download(name, lane)), "%s lane %d scan %d" % (name, lane, k)
    a.close(); b.close()


def test_step_argument_and_capacity_errors():
    sensor = "vlp16"
    seq = SynthSequence(sensor, seed=3)
    xyzi, ring = seq.scan(0)
    pl = Pipeline(make_params(sensor), lanes=1, max_points=1000)         # far fewer points than a scan has
    x = np.zeros((1, len(xyzi), 4), np.float32); x[0] = xyzi
    r = ring.reshape(1, -1).copy()
    with pytest.raises(LegoError, match=r"\(-4\).*max_points"):
        pl.step(x, r, [len(xyzi)], [0.0])
    with pytest.raises(LegoError, match=r"\(-4\)"):
        pl.step(np.ascontiguousarray(x[:, :500]), np.ascontiguousarray(r[:, :500]), [600], [0.0])     # npts > stride_points
    with pytest.raises(LegoError, match=r"\(-4\)"):
        pl.step(np.ascontiguousarray(x[:, :500]), np.ascontiguousarray(r[:, :500]), [-1], [0.0])
    with pytest.raises(LegoError, match=r"\(-1\)"):
        pl.step_lanes(np.array([x.ctypes.data], np.uint64), np.array([r.ctypes.data], np.uint64), [500], [0.0], floats_per_point=5)
    with pytest.raises(LegoError, match=r"\(-1\).*null lane pointer"):
        pl.step_lanes(np.array([0], np.uint64), np.array([r.ctypes.data], np.uint64), [500], [0.0], floats_per_point=4)
    # named-array upload: bounds are checked against the array's size
    lib = load()
    buf = np.zeros(64, np.float32)
    assert lib.lego_pipeline_upload(pl.h, 0, b"transformCur", buf.ctypes.data, 7) == -1
    assert lib.lego_pipeline_upload(pl.h, 0, b"transformCur", buf.ctypes.data, -1) == -1
    assert lib.lego_pipeline_upload(pl.h, 0, b"transformCur", buf.ctypes.data, 6) == 6
    pl.close()


def test_keyframe_pool_overflow_is_reported():
    """lego_params.maxKeyFrames = 3: the fourth key frame cannot be saved; the step that learns of it (one mapping call
    later) and every later step return LEGO_ERR_CAPACITY naming the lane -- the pipeline does not silently map against a
    frozen key-frame list."""
    sensor = "vlp16"
    seq = SynthSequence(sensor, seed=11, scene="street", speed=4.0)
    pl = Pipeline(make_params(sensor, max_keyframes=3), lanes=2, max_points=seq.cap)
    failed_at = None
    for k in range(40):
        xyzi, ring = seq.scan(k)
        x = np.zeros((2, seq.cap, 4), np.float32); x[:, :len(xyzi)] = xyzi
        r = np.zeros((2, seq.cap), np.uint16); r[:, :len(ring)] = ring
        try:
            pl.step(x, r, [len(xyzi)] * 2, [k * 0.1] * 2)
        except LegoError as ex:
            assert "(-4)" in str(ex) and "lane 0" in str(ex) and "capacity" in str(ex)
            failed_at = failed_at if failed_at is not None else k
    # a key frame every mapping call at 4 m/s (scans 2, 6, 10, 14): the 4th is refused at scan 14, reported by scan 18
    assert failed_at is not None and failed_at <= 22, failed_at
    assert int(pl.download("moState")[1]) == 3
    pl.close()


def test_cuda_graph_replay_changes_nothing(monkeypatch):
    """Below 148 lanes the per-scan launch sequences (front end, mapping body) are replayed from CUDA graphs
    (csrc/engine.cu: runGraphed; LEGO_GRAPH=0 keeps plain launches).  Same scans through both: images, features, poses,
    LM state and the key-frame count must be bit-identical, the graph handle must really have replayed graphs and both
    must have run the same number of kernels."""
    sensor = "vlp16"
    seqs = [SynthSequence(sensor, seed=41, scene="street", speed=4.0), SynthSequence(sensor, seed=42, scene="street", speed=9.0, dropout=0.1)]
    cap = seqs[0].cap
    monkeypatch.setenv("LEGO_GRAPH", "0")
    eager = Pipeline(make_params(sensor), lanes=2, max_points=cap)
    monkeypatch.setenv("LEGO_GRAPH", "1")
    graph = Pipeline(make_params(sensor), lanes=2, max_points=cap)
    mapped = 0
    for k in range(64):
        x = np.zeros((2, cap, 4), np.float32); r = np.zeros((2, cap), np.uint16); n = []
        for l, q in enumerate(seqs):
            xyzi, ring = q.scan(k)
            x[l, :len(xyzi)] = xyzi; r[l, :len(ring)] = ring; n.append(len(xyzi))
        out = []
        for pl in (eager, graph):
            pl.step(x, r, n, [seqs[0].stamp(k)] * 2, stage_mask=7)
            poses, flags = pl.poses()
            rec = [poses.copy(), flags.copy()]
            if k % 8 == 7 or k < 4:
                for l in range(2):
                    rec += [pl.download(nm, l).copy() for nm in NAMES]
                    rec += [pl.download("moState", l).copy(), pl.download("transformTobeMapped", l).copy()]
            out.append(rec)
        assert len(out[0]) == len(out[1])
        for a, b in zip(out[0], out[1]):
            assert _same(np.asarray(a), np.asarray(b)), "scan %d: graph replay differs from plain launches" % k
        mapped += int(np.count_nonzero(out[0][1] & 8))
    assert mapped >= 20
    assert eager.graph_launch_count() == 0
    # 10 front-end keys (5 input slots x 2 parities) seen twice before the first replay; mapping graphs on top
    assert graph.graph_launch_count() >= 64 - 20
    assert graph.launch_count() == eager.launch_count()
    eager.close(); graph.close()
