"""Host pose graph of the product (lego_pg_*, csrc/pose_graph.h: closed-form Jacobians, envelope Cholesky) against the
oracle's Gauss-Newton (oracle/thirdparty/tp_pose_graph.h: central-difference Jacobians).  No GPU involved: the graph is
host code in the reference too (GTSAM iSAM2, mapOptmization.cpp:1375-1409, 929-942).  GTSAM itself is not vendored in
the reference, so this pins the product against an independent formulation, not against GTSAM: parity unpinned."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))

import lego_loam_b200  # noqa: F401,E402
from lego_loam_b200 import capi  # noqa: E402
import refpy  # noqa: E402

VAR = [1e-6, 1e-6, 1e-6, 1e-8, 1e-8, 1e-6]   # mapOptmization.cpp:346-350


def _circle(n, rng, drift):
    """key poses on a circle (rx, ry, rz, tx, ty, tz) with accumulated drift, like a mapped loop before closure"""
    poses = []
    for k in range(n):
        a = 2 * np.pi * k / n
        poses.append([0.01 * np.sin(3 * a) + drift * k * 0.3, 0.02 * np.cos(2 * a), a * (1 + drift),
                      4 * np.sin(a) + drift * k, -4 * np.cos(a) - 0.5 * drift * k, 0.05 * np.sin(a) + 0.2 * drift * k])
    return np.array(poses) + rng.normal(0, 1e-4, (n, 6))


def _run(n, loops, steps, seed):
    rng = np.random.default_rng(seed)
    poses = _circle(n, rng, 2e-3)
    g = capi.PoseGraph()
    factors, meas, var = [], [], []
    g.add_prior(0, poses[0], VAR); g.insert(0, poses[0])
    factors.append([-1, 0]); meas.append(list(poses[0]) + [0] * 6); var.append(VAR)
    for k in range(1, n):
        g.add_between(k - 1, k, poses[k - 1], poses[k], VAR); g.insert(k, poses[k])
        factors.append([k - 1, k]); meas.append(list(poses[k - 1]) + list(poses[k])); var.append(VAR)
    for (a, b, noise) in loops:
        # the loop says: key a sits where key b is, up to a small offset (what ICP measures)
        frm = poses[b] + np.array([0, 0, 0, 0.02, -0.01, 0.005]); to = poses[b]
        g.add_between(a, b, frm, to, [noise] * 6)
        factors.append([a, b]); meas.append(list(frm) + list(to)); var.append([noise] * 6)
    for _ in range(steps):
        g.update()
    got = g.estimate()
    want = refpy.pose_graph(poses, factors, meas, var, steps)
    g.close()
    return poses, got, want


def test_chain_only_is_pass_through():
    poses, got, want = _run(40, [], 3, 1)
    # translation exactly, angles through the Rot3 round trip
    assert np.array_equal(got[:, :3], poses[:, 3:])
    d = got[:, 3:] - poses[:, :3]
    assert np.max(np.abs((d + np.pi) % (2 * np.pi) - np.pi)) < 1e-12


def test_one_loop_matches_oracle_every_step():
    for steps in (1, 2, 4, 8):
        poses, got, want = _run(80, [(79, 2, 0.03)], steps, 2)
        assert np.max(np.abs(got - want)) < 1e-7, steps
    # the loop moved the graph (it is not a no-op) and the prior holds key 0
    assert np.max(np.abs(got[:, :3] - poses[:, 3:])) > 1e-3
    assert np.max(np.abs(got[0, :3] - poses[0, 3:])) < 1e-5


def test_several_loops_converge_to_the_same_optimum():
    poses, got, want = _run(120, [(60, 3, 0.05), (100, 40, 0.02), (119, 1, 0.03)], 12, 3)
    assert np.max(np.abs(got - want)) < 1e-7
