// lo_io.hpp — data formats either side of the registration path, header-only, no CUDA.
//
// The reference takes its scans from a ROS 2 topic (pcl::fromROSMsg, src/lidar_odometry_node.cpp:47-48, point layout
// src/lidar_point_type.h:13-31) and its only file input is the PCD read of its test (pcl::io::loadPCDFile,
// test/test.cpp:194; fixture test/test_data/intersection00056.pcd: PCD v0.7, DATA binary, padded fields). A
// headless replacement needs the same records from files:
//   * PCD v0.7 reader (ascii / binary / binary_compressed, any field order, `_` padding fields, F/I/U of size 1-8)
//     into the three wire structs of include/lo_b200.h, and a writer (ascii / binary) for them;
//   * raw scan files: N x 32-byte PointXYZIRT records, exactly what lo_odometry_process takes;
//   * trajectory output: TUM (`t tx ty tz qx qy qz qw`) and KITTI (row-major 3x4) text, one line per scan — what the
//     node publishes as TF + nav_msgs/Odometry (src/lidar_odometry_node.cpp:76-107), as a file.
// Errors are reported by throwing lo::io::Error (std::runtime_error).
#pragma once

#include <cstdint>
#include <cstdio>
#include <cstring>
#include <fstream>
#include <sstream>
#include <stdexcept>
#include <string>
#include <vector>

#include "lo_b200.hpp"

namespace lo {
namespace io {

struct Error : std::runtime_error {
    using std::runtime_error::runtime_error;
};

// ---- PCD ----------------------------------------------------------------------------------------------------
struct PcdField {
    std::string name;
    uint32_t size = 4, count = 1, offset = 0;   // offset into the record of DATA binary
    char type = 'F';
};
struct PcdHeader {
    std::vector<PcdField> fields;
    uint64_t width = 0, height = 1, points = 0;
    uint32_t record_bytes = 0;
    std::string data;                 // "ascii" | "binary" | "binary_compressed"
    size_t data_offset = 0;           // first byte after the DATA line
    const PcdField* find(const std::string& name) const {
        for (const auto& f : fields)
            if (f.name == name) return &f;
        return nullptr;
    }
};

inline std::string read_file(const std::string& path) {
    std::ifstream in(path, std::ios::binary);
    if (!in) throw Error("cannot open " + path);
    std::ostringstream ss;
    ss << in.rdbuf();
    return ss.str();
}

inline PcdHeader parse_pcd_header(const std::string& blob) {
    PcdHeader h;
    size_t pos = 0;
    bool have_points = false;
    std::vector<std::string> names, sizes, types, counts;
    auto split = [](const std::string& line) {
        std::vector<std::string> out;
        std::istringstream ss(line);
        for (std::string w; ss >> w;) out.push_back(w);
        return out;
    };
    while (pos < blob.size()) {
        size_t eol = blob.find('\n', pos);
        if (eol == std::string::npos) eol = blob.size();
        std::string line = blob.substr(pos, eol - pos);
        pos = eol + 1;
        if (!line.empty() && line.back() == '\r') line.pop_back();
        if (line.empty() || line[0] == '#') continue;
        auto w = split(line);
        if (w.empty()) continue;
        const std::string key = w[0];
        w.erase(w.begin());
        if (key == "FIELDS" || key == "COLUMNS") names = w;
        else if (key == "SIZE") sizes = w;
        else if (key == "TYPE") types = w;
        else if (key == "COUNT") counts = w;
        else if (key == "WIDTH" && !w.empty()) h.width = std::stoull(w[0]);
        else if (key == "HEIGHT" && !w.empty()) h.height = std::stoull(w[0]);
        else if (key == "POINTS" && !w.empty()) { h.points = std::stoull(w[0]); have_points = true; }
        else if (key == "DATA") {
            if (w.empty()) throw Error("PCD: DATA line without a kind");
            h.data = w[0];
            h.data_offset = pos;
            break;
        }
    }
    if (h.data.empty()) throw Error("PCD: no DATA line");
    if (names.empty() || sizes.size() != names.size() || types.size() != names.size()) throw Error("PCD: FIELDS / SIZE / TYPE disagree");
    if (!counts.empty() && counts.size() != names.size()) throw Error("PCD: COUNT disagrees with FIELDS");
    uint32_t off = 0;
    for (size_t i = 0; i < names.size(); i++) {
        PcdField f;
        f.name = names[i];
        f.size = static_cast<uint32_t>(std::stoul(sizes[i]));
        f.type = types[i].empty() ? 'F' : types[i][0];
        f.count = counts.empty() ? 1u : static_cast<uint32_t>(std::stoul(counts[i]));
        f.offset = off;
        if (f.size == 0 || f.size > 8 || (f.type != 'F' && f.type != 'I' && f.type != 'U')) throw Error("PCD: unsupported field type");
        off += f.size * f.count;
        h.fields.push_back(f);
    }
    h.record_bytes = off;
    if (!have_points) h.points = h.width * h.height;
    return h;
}

// LZF (the codec of DATA binary_compressed): literal runs `000LLLLL` (L+1 bytes) and back references
// `LLLooooo oooooooo` (length L+2, or 7 + next byte + 2) to `offset + 1` bytes back.
inline std::string lzf_decompress(const unsigned char* in, size_t n_in, size_t n_out) {
    std::string out(n_out, '\0');
    size_t ip = 0, op = 0;
    while (ip < n_in) {
        const unsigned ctrl = in[ip++];
        if (ctrl < 32) {
            const size_t run = ctrl + 1;
            if (ip + run > n_in || op + run > n_out) throw Error("PCD: corrupt compressed data");
            std::memcpy(&out[op], in + ip, run);
            ip += run;
            op += run;
        } else {
            size_t len = ctrl >> 5;
            if (len == 7) {
                if (ip >= n_in) throw Error("PCD: corrupt compressed data");
                len += in[ip++];
            }
            if (ip >= n_in) throw Error("PCD: corrupt compressed data");
            const size_t back = ((ctrl & 0x1Fu) << 8 | in[ip++]) + 1;
            len += 2;
            if (back > op || op + len > n_out) throw Error("PCD: corrupt compressed data");
            for (size_t k = 0; k < len; k++, op++) out[op] = out[op - back];   // may overlap: byte by byte
        }
    }
    if (op != n_out) throw Error("PCD: compressed data is shorter than announced");
    return out;
}

namespace detail {
inline double load_scalar(const unsigned char* p, const PcdField& f) {
    switch (f.type) {
        case 'F':
            if (f.size == 4) { float v; std::memcpy(&v, p, 4); return v; }
            if (f.size == 8) { double v; std::memcpy(&v, p, 8); return v; }
            break;
        case 'I':
            if (f.size == 1) { int8_t v; std::memcpy(&v, p, 1); return v; }
            if (f.size == 2) { int16_t v; std::memcpy(&v, p, 2); return v; }
            if (f.size == 4) { int32_t v; std::memcpy(&v, p, 4); return v; }
            if (f.size == 8) { int64_t v; std::memcpy(&v, p, 8); return static_cast<double>(v); }
            break;
        case 'U':
            if (f.size == 1) { uint8_t v; std::memcpy(&v, p, 1); return v; }
            if (f.size == 2) { uint16_t v; std::memcpy(&v, p, 2); return v; }
            if (f.size == 4) { uint32_t v; std::memcpy(&v, p, 4); return v; }
            if (f.size == 8) { uint64_t v; std::memcpy(&v, p, 8); return static_cast<double>(v); }
            break;
    }
    throw Error("PCD: unsupported field size");
}

// a column of the file as doubles (first element of multi-count fields); empty if the field is absent
struct Column {
    std::vector<double> v;
    bool present() const { return !v.empty(); }
};

struct Table {   // the file decoded into per-field columns on demand
    PcdHeader h;
    std::string blob, soa;   // raw file; de-compressed column-major payload (binary_compressed only)
    std::vector<std::vector<double>> ascii;   // ascii: one vector per scalar column

    explicit Table(const std::string& path) : blob(read_file(path)) {
        h = parse_pcd_header(blob);
        const unsigned char* d = reinterpret_cast<const unsigned char*>(blob.data()) + h.data_offset;
        const size_t left = blob.size() - h.data_offset;
        if (h.data == "binary") {
            if (left < h.points * h.record_bytes) throw Error("PCD: file is shorter than POINTS x record size");
        } else if (h.data == "binary_compressed") {
            if (left < 8) throw Error("PCD: truncated compressed header");
            uint32_t n_comp, n_raw;
            std::memcpy(&n_comp, d, 4);
            std::memcpy(&n_raw, d + 4, 4);
            if (left < 8 + static_cast<size_t>(n_comp) || n_raw != h.points * h.record_bytes) throw Error("PCD: compressed sizes disagree with the header");
            soa = lzf_decompress(d + 8, n_comp, n_raw);
        } else if (h.data == "ascii") {
            size_t cols = 0;
            for (const auto& f : h.fields) cols += f.count;
            ascii.assign(cols, std::vector<double>());
            for (auto& c : ascii) c.reserve(h.points);
            const char* p = blob.data() + h.data_offset;
            const char* end = blob.data() + blob.size();
            for (uint64_t i = 0; i < h.points; i++)
                for (size_t c = 0; c < cols; c++) {
                    while (p < end && (*p == ' ' || *p == '\n' || *p == '\r' || *p == '\t')) p++;
                    if (p >= end) throw Error("PCD: ascii data ends early");
                    char* stop = nullptr;
                    ascii[c].push_back(std::strtod(p, &stop));
                    if (stop == p) throw Error("PCD: unreadable ascii value");
                    p = stop;
                }
        } else {
            throw Error("PCD: unknown DATA kind " + h.data);
        }
    }

    Column column(const std::string& name) const {
        Column out;
        const PcdField* f = h.find(name);
        if (!f) return out;
        out.v.resize(h.points);
        if (h.data == "ascii") {
            size_t c = 0;
            for (const auto& g : h.fields) {
                if (&g == f) break;
                c += g.count;
            }
            out.v = ascii[c];
        } else if (h.data == "binary") {
            const unsigned char* d = reinterpret_cast<const unsigned char*>(blob.data()) + h.data_offset + f->offset;
            for (uint64_t i = 0; i < h.points; i++) out.v[i] = load_scalar(d + i * h.record_bytes, *f);
        } else {   // binary_compressed stores field after field
            const unsigned char* d = reinterpret_cast<const unsigned char*>(soa.data()) + static_cast<size_t>(f->offset) * h.points;
            for (uint64_t i = 0; i < h.points; i++) out.v[i] = load_scalar(d + i * static_cast<size_t>(f->size) * f->count, *f);
        }
        return out;
    }
    Column first_of(std::initializer_list<const char*> names) const {
        for (const char* n : names) {
            Column c = column(n);
            if (c.present() || (h.find(n) && h.points == 0)) return c;
        }
        return Column();
    }
};
}  // namespace detail

// x y z  ->  pcl::PointXYZ (pcl::io::loadPCDFile<pcl::PointXYZ>, test/test.cpp:194)
inline void read_pcd(const std::string& path, pcl::PointCloud<pcl::PointXYZ>& cloud) {
    const detail::Table t(path);
    const auto x = t.column("x"), y = t.column("y"), z = t.column("z");
    if (t.h.points && !(x.present() && y.present() && z.present())) throw Error("PCD: " + path + " has no x/y/z fields");
    cloud.points.assign(t.h.points, pcl::PointXYZ());
    for (uint64_t i = 0; i < t.h.points; i++) cloud.points[i] = pcl::PointXYZ(static_cast<float>(x.v[i]), static_cast<float>(y.v[i]), static_cast<float>(z.v[i]));
    cloud.width = static_cast<uint32_t>(t.h.width);
    cloud.height = static_cast<uint32_t>(t.h.height);
}

// x y z normal_x normal_y normal_z [curvature]  ->  pcl::PointNormal
inline void read_pcd(const std::string& path, pcl::PointCloud<pcl::PointNormal>& cloud) {
    const detail::Table t(path);
    const auto x = t.column("x"), y = t.column("y"), z = t.column("z");
    const auto nx = t.column("normal_x"), ny = t.column("normal_y"), nz = t.column("normal_z"), cv = t.column("curvature");
    if (t.h.points && !(x.present() && y.present() && z.present())) throw Error("PCD: " + path + " has no x/y/z fields");
    cloud.points.assign(t.h.points, pcl::PointNormal());
    for (uint64_t i = 0; i < t.h.points; i++) {
        pcl::PointNormal p(static_cast<float>(x.v[i]), static_cast<float>(y.v[i]), static_cast<float>(z.v[i]));
        if (nx.present() && ny.present() && nz.present()) {
            p.normal_x = static_cast<float>(nx.v[i]);
            p.normal_y = static_cast<float>(ny.v[i]);
            p.normal_z = static_cast<float>(nz.v[i]);
        }
        if (cv.present()) p.curvature = static_cast<float>(cv.v[i]);
        cloud.points[i] = p;
    }
    cloud.width = static_cast<uint32_t>(t.h.width);
    cloud.height = static_cast<uint32_t>(t.h.height);
}

// x y z intensity ring time  ->  lidar_point::PointXYZIRT (the registered fields of src/lidar_point_type.h:24-31;
// `t` / `timestamp` are accepted for `time`, as other drivers name it)
inline void read_pcd(const std::string& path, pcl::PointCloud<lidar_point::PointXYZIRT>& cloud) {
    const detail::Table t(path);
    const auto x = t.column("x"), y = t.column("y"), z = t.column("z");
    const auto in = t.column("intensity"), ring = t.column("ring"), tm = t.first_of({"time", "t", "timestamp"});
    if (t.h.points && !(x.present() && y.present() && z.present())) throw Error("PCD: " + path + " has no x/y/z fields");
    if (t.h.points && !(ring.present() && tm.present())) throw Error("PCD: " + path + " has no ring / time fields (needed by the classifier and the deskew)");
    cloud.points.assign(t.h.points, lidar_point::PointXYZIRT());
    for (uint64_t i = 0; i < t.h.points; i++) {
        lidar_point::PointXYZIRT p;
        p.x = static_cast<float>(x.v[i]);
        p.y = static_cast<float>(y.v[i]);
        p.z = static_cast<float>(z.v[i]);
        if (in.present()) p.intensity = static_cast<float>(in.v[i]);
        p.ring = static_cast<uint16_t>(ring.v[i]);
        p.time = static_cast<float>(tm.v[i]);
        cloud.points[i] = p;
    }
    cloud.width = static_cast<uint32_t>(t.h.width);
    cloud.height = static_cast<uint32_t>(t.h.height);
}

namespace detail {
inline void write_header(std::ostream& out, const char* fields, const char* sizes, const char* types, const char* counts, size_t n, bool binary) {
    out << "# .PCD v0.7 - Point Cloud Data file format\nVERSION 0.7\nFIELDS " << fields << "\nSIZE " << sizes << "\nTYPE " << types << "\nCOUNT "
        << counts << "\nWIDTH " << n << "\nHEIGHT 1\nVIEWPOINT 0 0 0 1 0 0 0\nPOINTS " << n << "\nDATA " << (binary ? "binary" : "ascii") << "\n";
}
inline std::ofstream open_out(const std::string& path) {
    std::ofstream out(path, std::ios::binary);
    if (!out) throw Error("cannot create " + path);
    out.precision(9);
    return out;
}
}  // namespace detail

// binary files carry the in-memory records unchanged (padding as `_` fields), the way PCL writes them
inline void write_pcd(const std::string& path, const pcl::PointCloud<pcl::PointXYZ>& cloud, bool binary = true) {
    auto out = detail::open_out(path);
    if (binary) {
        detail::write_header(out, "x y z _", "4 4 4 4", "F F F U", "1 1 1 1", cloud.size(), true);
        out.write(reinterpret_cast<const char*>(cloud.points.data()), static_cast<std::streamsize>(cloud.size() * sizeof(pcl::PointXYZ)));
    } else {
        detail::write_header(out, "x y z", "4 4 4", "F F F", "1 1 1", cloud.size(), false);
        for (const auto& p : cloud.points) out << p.x << ' ' << p.y << ' ' << p.z << '\n';
    }
    if (!out) throw Error("write failed: " + path);
}

inline void write_pcd(const std::string& path, const pcl::PointCloud<pcl::PointNormal>& cloud, bool binary = true) {
    auto out = detail::open_out(path);
    if (binary) {
        detail::write_header(out, "x y z _ normal_x normal_y normal_z _ curvature _", "4 4 4 4 4 4 4 4 4 4", "F F F U F F F U F U", "1 1 1 1 1 1 1 1 1 3",
                             cloud.size(), true);
        out.write(reinterpret_cast<const char*>(cloud.points.data()), static_cast<std::streamsize>(cloud.size() * sizeof(pcl::PointNormal)));
    } else {
        detail::write_header(out, "x y z normal_x normal_y normal_z curvature", "4 4 4 4 4 4 4", "F F F F F F F", "1 1 1 1 1 1 1", cloud.size(), false);
        for (const auto& p : cloud.points)
            out << p.x << ' ' << p.y << ' ' << p.z << ' ' << p.normal_x << ' ' << p.normal_y << ' ' << p.normal_z << ' ' << p.curvature << '\n';
    }
    if (!out) throw Error("write failed: " + path);
}

inline void write_pcd(const std::string& path, const pcl::PointCloud<lidar_point::PointXYZIRT>& cloud, bool binary = true) {
    auto out = detail::open_out(path);
    if (binary) {
        detail::write_header(out, "x y z _ intensity ring _ time _", "4 4 4 4 4 2 2 4 4", "F F F U F U U F U", "1 1 1 1 1 1 1 1 1", cloud.size(), true);
        out.write(reinterpret_cast<const char*>(cloud.points.data()), static_cast<std::streamsize>(cloud.size() * sizeof(lidar_point::PointXYZIRT)));
    } else {
        detail::write_header(out, "x y z intensity ring time", "4 4 4 4 2 4", "F F F F U F", "1 1 1 1 1 1", cloud.size(), false);
        for (const auto& p : cloud.points) out << p.x << ' ' << p.y << ' ' << p.z << ' ' << p.intensity << ' ' << p.ring << ' ' << p.time << '\n';
    }
    if (!out) throw Error("write failed: " + path);
}

// ---- raw scans: N x 32-byte PointXYZIRT, the record lo_odometry_process takes -----------------------------------
inline void read_scan_bin(const std::string& path, pcl::PointCloud<lidar_point::PointXYZIRT>& cloud) {
    const std::string blob = read_file(path);
    if (blob.size() % sizeof(lidar_point::PointXYZIRT) != 0) throw Error(path + ": size is not a multiple of the 32-byte PointXYZIRT record");
    cloud.points.resize(blob.size() / sizeof(lidar_point::PointXYZIRT));
    if (!blob.empty()) std::memcpy(static_cast<void*>(cloud.points.data()), blob.data(), blob.size());
    cloud.width = static_cast<uint32_t>(cloud.points.size());
    cloud.height = 1;
}

inline void write_scan_bin(const std::string& path, const pcl::PointCloud<lidar_point::PointXYZIRT>& cloud) {
    auto out = detail::open_out(path);
    out.write(reinterpret_cast<const char*>(cloud.points.data()), static_cast<std::streamsize>(cloud.size() * sizeof(lidar_point::PointXYZIRT)));
    if (!out) throw Error("write failed: " + path);
}

// by extension: .pcd or anything else = raw records
inline void read_scan(const std::string& path, pcl::PointCloud<lidar_point::PointXYZIRT>& cloud) {
    if (path.size() >= 4 && path.compare(path.size() - 4, 4, ".pcd") == 0) read_pcd(path, cloud);
    else read_scan_bin(path, cloud);
}

// ---- trajectory -------------------------------------------------------------------------------------------
class TrajectoryWriter {
public:
    enum Format { TUM, KITTI };
    TrajectoryWriter(const std::string& path, Format format) : out_(detail::open_out(path)), format_(format) {}
    // TUM: `timestamp tx ty tz qx qy qz qw`; KITTI: the 3x4 [R|t], row-major
    void add(double timestamp, const Pose3D& pose) {
        const auto& t = pose.translation;
        const auto& q = pose.rotation;
        if (format_ == TUM) {
            char line[256];
            std::snprintf(line, sizeof(line), "%.6f %.9g %.9g %.9g %.9g %.9g %.9g %.9g\n", timestamp, t.x(), t.y(), t.z(), q.x(), q.y(), q.z(), q.w());
            out_ << line;
        } else {
            float r[9];
            pose.rotationMatrix(r);
            char line[512];
            std::snprintf(line, sizeof(line), "%.9g %.9g %.9g %.9g %.9g %.9g %.9g %.9g %.9g %.9g %.9g %.9g\n", r[0], r[1], r[2], t.x(), r[3], r[4], r[5],
                          t.y(), r[6], r[7], r[8], t.z());
            out_ << line;
        }
        if (!out_) throw Error("trajectory write failed");
    }
    void flush() { out_.flush(); }

private:
    std::ofstream out_;
    Format format_;
};

}  // namespace io
}  // namespace lo
