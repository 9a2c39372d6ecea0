#!/usr/bin/env python3
"""TEST INFRASTRUCTURE ONLY (oracle).

Writes oracle/_ref/cfg/<name>/utility.h = the reference's LeGO-LOAM/include/utility.h with ONLY the sensor
block (utility.h:63-68) and, where a config asks for it, the useCloudRing line (utility.h:60) replaced.  The
generated header is placed first on the include path so the reference's verbatim node sources are re-targeted
to another sensor without editing them (the reference changes sensor the same way: edit + rebuild, README.md:60-86).
Outputs land under oracle/_ref/ (git-ignored); reference sources are never copied into tracked files.
"""
import os
import sys

REF = os.environ.get("LEGO_REF_ROOT", "/root/reference")
HERE = os.path.dirname(os.path.abspath(__file__))

# name -> (N_SCAN, Horizon_SCAN, ang_res_x, ang_res_y, ang_bottom, groundScanInd, useCloudRing[, loopClosureEnableFlag])
CONFIGS = {
    # utility.h:63-68 as shipped
    "vlp16": ("16", "1800", "0.2", "2.0", "15.0+0.1", "7", "true"),
    # same sensor, projection by elevation angle instead of the ring channel (utility.h:60 = false)
    "vlp16_noring": ("16", "1800", "0.2", "2.0", "15.0+0.1", "7", "false"),
    # README.md:73-82 ("README parameters" of BASELINE config 3; N_Scan typo corrected to N_SCAN)
    "hdl32": ("32", "1800", "360.0/float(Horizon_SCAN)", "41.333/float(N_SCAN-1)", "30.666666", "20", "true"),
    # not in the reference (README.md:86): 64-beam 64x2083, parameters from SURVEY.md section 8c
    "hdl64": ("64", "2083", "360.0/float(Horizon_SCAN)", "26.9/float(N_SCAN-1)", "24.9", "50", "true"),
    # utility.h:79-84
    "vls128": ("128", "1800", "0.2", "0.3", "25.0", "10", "true"),
    # utility.h:88-102 (Ouster OS1-16 / OS1-64: 1024-column images)
    "os1_16": ("16", "1024", "360.0/float(Horizon_SCAN)", "33.2/float(N_SCAN-1)", "16.6+0.1", "7", "true"),
    "os1_64": ("64", "1024", "360.0/float(Horizon_SCAN)", "33.2/float(N_SCAN-1)", "16.6+0.1", "15", "true"),
    # VLP-16 with loop closure switched on (utility.h:104 = true), SURVEY.md section 8f n2 / BASELINE config 5
    "vlp16_loop": ("16", "1800", "0.2", "2.0", "15.0+0.1", "7", "true", "true"),
}


def generate(name, out_root):
    n_scan, hz, rx, ry, bottom, gsi, ring = CONFIGS[name][:7]
    loop = CONFIGS[name][7] if len(CONFIGS[name]) > 7 else "false"
    src = open(os.path.join(REF, "LeGO-LOAM/include/utility.h")).read().split("\n")
    # lines are 1-based in citations: 60 = useCloudRing, 63..68 = VLP-16 block
    assert "useCloudRing" in src[59] and "N_SCAN = 16" in src[62] and "groundScanInd" in src[67], "utility.h layout changed"
    src[59] = "extern const bool useCloudRing = %s;" % ring
    src[62] = "extern const int N_SCAN = %s;" % n_scan
    src[63] = "extern const int Horizon_SCAN = %s;" % hz
    src[64] = "extern const float ang_res_x = %s;" % rx
    src[65] = "extern const float ang_res_y = %s;" % ry
    src[66] = "extern const float ang_bottom = %s;" % bottom
    src[67] = "extern const int groundScanInd = %s;" % gsi
    assert "loopClosureEnableFlag" in src[103], "utility.h layout changed"
    src[103] = "extern const bool loopClosureEnableFlag = %s;" % loop
    d = os.path.join(out_root, "cfg", name)
    os.makedirs(d, exist_ok=True)
    with open(os.path.join(d, "utility.h"), "w") as f:
        f.write("\n".join(src))
    return d


if __name__ == "__main__":
    out_root = os.path.join(HERE, "_ref")
    for nm in (sys.argv[1:] or list(CONFIGS)):
        print(generate(nm, out_root))
