// TEST INFRASTRUCTURE (oracle) -- restatement of pcl::io::savePCDFileASCII for a pcl::PointCloud<pcl::PointXYZI>.
// PCL (>= 1.7, the version ROS kinetic/melodic ship) is a third-party dependency of the reference and is not vendored
// under /root/reference: PARITY UNPINNED for the byte layout.  What is restated, from the published writer
// (pcl/io/impl/pcd_io.hpp, PCDWriter::writeASCII<PointT> + PCDWriter::generateHeader<PointT>):
//   - header lines "# .PCD v0.7 - Point Cloud Data file format", VERSION 0.7, FIELDS / SIZE / TYPE / COUNT of the
//     real fields only (the padding fields named "_" are skipped), WIDTH = points, HEIGHT = 1 for an unorganised
//     cloud, VIEWPOINT 0 0 0 1 0 0 0 (sensor origin / identity orientation), POINTS, DATA ascii;
//   - one line per point, fields separated by one blank, values written through an ostream in the classic locale with
//     precision 8 (the default of savePCDFileASCII), NaN as "nan", the line trimmed.
// Call site in the reference: mapOptmization.cpp:731, 753-755.
#pragma once
#include <cmath>
#include <cstdio>
#include <string>
#include <vector>

namespace tp {

// An ostream in the classic locale with precision 8 and default float flags formats a float the way "%.8g" formats
// the same value widened to double ([facet.num.put.virtuals]); written with snprintf so that this checker shares no
// formatting code with the product's writer.
inline std::string pcd_ascii_xyzi(const std::vector<float>& xyzi) {
    const size_t n = xyzi.size() / 4;
    char buf[64];
    std::string out;
    out.reserve(256 + n * 48);
    out += "# .PCD v0.7 - Point Cloud Data file format\n";
    out += "VERSION 0.7\n";
    out += "FIELDS x y z intensity\n";
    out += "SIZE 4 4 4 4\n";
    out += "TYPE F F F F\n";
    out += "COUNT 1 1 1 1\n";
    std::snprintf(buf, sizeof(buf), "WIDTH %zu\nHEIGHT 1\n", n); out += buf;
    out += "VIEWPOINT 0 0 0 1 0 0 0\n";
    std::snprintf(buf, sizeof(buf), "POINTS %zu\n", n); out += buf;
    out += "DATA ascii\n";
    for (size_t i = 0; i < n; ++i) {
        for (int d = 0; d < 4; ++d) {
            const float value = xyzi[4 * i + d];
            if (std::isnan(value)) out += "nan";
            else { std::snprintf(buf, sizeof(buf), "%.8g", static_cast<double>(value)); out += buf; }
            out += d < 3 ? " " : "\n";
        }
    }
    return out;
}

}  // namespace tp
