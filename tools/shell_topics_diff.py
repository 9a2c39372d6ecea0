"""Debug aid for tests/test_gpu_ros_shells.py: feeds the same scans to the reference's nodes and to the B200 node shells and lists
EVERY topic that differs (the test stops at the first)."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "oracle"))
import lego_loam_b200  # noqa: E402,F401
import refpy  # noqa: E402
from lego_loam_b200.synth import SynthSequence  # noqa: E402

CLOUDS = ["/segmented_cloud", "/outlier_cloud", "/laser_cloud_sharp", "/laser_cloud_less_sharp", "/laser_cloud_flat",
          "/laser_cloud_less_flat", "/laser_cloud_corner_last", "/laser_cloud_surf_last", "/outlier_cloud_last"]
n = int(sys.argv[1]) if len(sys.argv) > 1 else 30
seq = SynthSequence("vlp16", seed=21, scene="street", speed=4.0)
ref = refpy.RefPipeline("vlp16", snapshots=False)
ref.set_verbatim(True)
refpy.ref_subscribe_all(ref, True)
sh = refpy.ShellPipeline("vlp16", subscribe_all=True)
bad = 0
for k in range(n):
    xyzi, ring = seq.scan(k)
    rf = ref.feed(xyzi, ring, seq.stamp(k), stages=15)
    sf = sh.feed(xyzi, ring, seq.stamp(k))
    msgs = []
    if sf < 0 or sh.errors():
        msgs.append("shell error sf=%d errors=%d" % (sf, sh.errors()))
    if bool(sf & 8) != bool(rf & 8):
        msgs.append("mapping flag %d vs %d" % (sf, rf))
    for topic in CLOUDS + list(refpy.TOPICS_ODOMETRY):
        a, b = sh.topic(topic), refpy.ref_topic(ref, topic)
        if (a is None) != (b is None):
            msgs.append("%s: only %s published" % (topic, "shells" if b is None else "reference")); continue
        if a is None:
            continue
        if a["count"] != b["count"]:
            msgs.append("%s: count %d vs %d" % (topic, a["count"], b["count"]))
        hdr = lambda m: (m["frame_id"], m.get("child_frame_id", ""), m["stamp"])
        if hdr(a) != hdr(b):
            msgs.append("%s: header %r vs %r" % (topic, hdr(a), hdr(b)))
        if a["values"].shape != b["values"].shape:
            msgs.append("%s: %d vs %d values" % (topic, len(a["values"]), len(b["values"])))
        else:
            d = float(np.max(np.abs(a["values"] - b["values"]))) if len(a["values"]) else 0.0
            if d > 2e-4:
                msgs.append("%s: max abs diff %.3g" % (topic, d))
    print("scan %2d flags shells %d ref %d %s" % (k, sf, rf, "ok" if not msgs else ""))
    for m in msgs:
        print("    " + m)
    bad += len(msgs)
print("differences:", bad)
