"""CPU: lego_write_pcd_ascii (host file output, no device needed) writes, for the clouds the reference saves at shutdown
(mapOptmization.cpp:729-755), the bytes the oracle's statement of pcl::io::savePCDFileASCII produces
(oracle/thirdparty/tp_pcd.h; PCL is not vendored in the reference: the byte layout is parity-unpinned)."""
import numpy as np

import lego_loam_b200  # noqa: F401
from conftest import require_ref
from lego_loam_b200 import capi
from lego_loam_b200.synth import SynthSequence


def test_pcd_files_of_the_reference_shutdown(tmp_path):
    import refpy
    require_ref("vlp16")
    seq = SynthSequence("vlp16", seed=2, scene="street", speed=5.0)
    ref = refpy.RefPipeline("vlp16")
    for k in range(30):
        xyzi, ring = seq.scan(k)
        ref.feed(xyzi, ring, seq.stamp(k), stages=7)
    assert ref.global_map() == 4
    sizes = {}
    for name in ("finalCloud", "cornerMap", "surfaceMap", "trajectory"):
        cloud = ref.cloud("pcd:" + name)
        sizes[name] = len(cloud)
        path = tmp_path / (name + ".pcd")
        capi.write_pcd_ascii(str(path), cloud)
        assert path.read_bytes() == ref.pcd_text(name), name
    assert sizes["trajectory"] >= 4 and sizes["cornerMap"] > 100 and sizes["surfaceMap"] > 1000, sizes
    # publishGlobalMap's cloud is what finalCloud.pcd holds (mapOptmization.cpp:731)
    assert np.array_equal(ref.cloud("globalMapKeyFramesDS"), ref.cloud("pcd:finalCloud"))
    ref.close()


def test_pcd_number_formatting_and_errors(tmp_path):
    c = np.array([[1.5, 2, 3, 0], [0.1, 1e-7, 123456.789, np.nan], [-1 / 3, 1e10, 0, 7]], np.float32)
    path = tmp_path / "t.pcd"
    capi.write_pcd_ascii(str(path), c)
    lines = path.read_text().split("\n")
    assert lines[:11] == ["# .PCD v0.7 - Point Cloud Data file format", "VERSION 0.7", "FIELDS x y z intensity", "SIZE 4 4 4 4", "TYPE F F F F",
                          "COUNT 1 1 1 1", "WIDTH 3", "HEIGHT 1", "VIEWPOINT 0 0 0 1 0 0 0", "POINTS 3", "DATA ascii"]
    assert lines[11:] == ["1.5 2 3 0", "0.1 1e-07 123456.79 nan", "-0.33333334 1e+10 0 7", ""]
    capi.write_pcd_ascii(str(path), np.zeros((0, 4), np.float32))
    assert path.read_text().endswith("WIDTH 0\nHEIGHT 1\nVIEWPOINT 0 0 0 1 0 0 0\nPOINTS 0\nDATA ascii\n")
    try:
        capi.write_pcd_ascii(str(tmp_path / "no_such_dir" / "x.pcd"), c)
    except capi.LegoError:
        pass
    else:
        raise AssertionError("writing into a missing directory must fail")
