// Row F4: the text of a ParaView .vtu piece from flat arrays.
//
// The reference writes one ASCII VTK XML UnstructuredGrid file per layer (padne/paraview.py:234-298),
// one <Piece> per mesh (padne/paraview.py:209-231).  Every DataArray payload is
// " ".join(str(value) for value in values) (padne/paraview.py:91) over lists gathered by walking the
// half-edge objects vertex by vertex and face by face (padne/paraview.py:121-124, 140-145, 159-181).
// Here the same bytes are produced from the arrays the solve already holds (coordinates, triangles, the
// voltage slice of v): host code only -- the output is a file -- formatted in parallel chunks.
//
// str(float) is CPython's repr: the shortest digit string that round-trips (std::to_chars gives the same
// digits), printed in fixed notation when the decimal exponent e satisfies -4 <= e < 16 (always with a
// fractional part, "2.0"), otherwise as d[.ddd]e[+-]XX with at least two exponent digits; "nan", "inf",
// "-inf", "-0.0" as CPython prints them.  tests/test_vtu.py checks it against repr() on random bit patterns.
#include <algorithm>
#include <charconv>
#include <cmath>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <thread>
#include <vector>

#include "padne_b200.h"

namespace {

constexpr int F64_WIDTH = 25;   // "-1.7976931348623157e+308" is 24 characters, plus the separator
constexpr int I64_WIDTH = 21;   // "-9223372036854775808" is 20
constexpr int I32_WIDTH = 12;   // "-2147483648" is 11
constexpr int64_t SERIAL_BELOW = 1 << 15;

inline char *put_bytes(char *p, const char *s) {
    size_t n = std::strlen(s);
    std::memcpy(p, s, n);
    return p + n;
}

inline char *put_i64(char *p, int64_t v) { return std::to_chars(p, p + I64_WIDTH, v).ptr; }

// CPython float.__repr__
inline char *put_f64(char *p, double x) {
    if (std::isnan(x)) return put_bytes(p, "nan");
    if (std::isinf(x)) return put_bytes(p, x < 0 ? "-inf" : "inf");
    if (x == 0.0) return put_bytes(p, std::signbit(x) ? "-0.0" : "0.0");
    char sci[32];
    char *end = std::to_chars(sci, sci + sizeof sci, x, std::chars_format::scientific).ptr;
    const char *s = sci;
    if (*s == '-') *p++ = *s++;
    // s: d[.ddd]e[+-]XX[X]
    char digits[24];
    int nd = 0;
    digits[nd++] = *s++;
    if (*s == '.') {
        ++s;
        while (*s != 'e') digits[nd++] = *s++;
    }
    ++s;   // 'e'
    bool eneg = *s == '-';
    ++s;
    int e = 0;
    for (; s < end; ++s) e = e * 10 + (*s - '0');
    if (eneg) e = -e;
    if (e < -4 || e >= 16) {
        *p++ = digits[0];
        if (nd > 1) {
            *p++ = '.';
            std::memcpy(p, digits + 1, nd - 1);
            p += nd - 1;
        }
        *p++ = 'e';
        *p++ = e < 0 ? '-' : '+';
        int a = e < 0 ? -e : e;
        if (a >= 100) *p++ = char('0' + a / 100);
        *p++ = char('0' + (a / 10) % 10);
        *p++ = char('0' + a % 10);
        return p;
    }
    if (e < 0) {   // 0.000ddd
        *p++ = '0';
        *p++ = '.';
        for (int k = 0; k < -e - 1; ++k) *p++ = '0';
        std::memcpy(p, digits, nd);
        return p + nd;
    }
    int ip = e + 1;   // digits before the point
    if (nd <= ip) {
        std::memcpy(p, digits, nd);
        p += nd;
        for (int k = nd; k < ip; ++k) *p++ = '0';
        *p++ = '.';
        *p++ = '0';
        return p;
    }
    std::memcpy(p, digits, ip);
    p += ip;
    *p++ = '.';
    std::memcpy(p, digits + ip, nd - ip);
    return p + (nd - ip);
}

int io_threads() {
    if (const char *s = std::getenv("PADNE_B200_IO_THREADS")) {
        int t = std::atoi(s);
        if (t >= 1) return std::min(t, 64);
    }
    unsigned hw = std::thread::hardware_concurrency();
    return int(std::clamp(hw, 1u, 16u));
}

// Items 0..n-1, each written by item(q, i) -> end and followed by one space; the last space is dropped.
// Parallel form: worker k formats its contiguous range into its own slot of the caller's buffer (slots are
// `width` bytes per item apart, so they cannot overlap), then the slots are closed up front to back.
template <class Item>
char *join_items(char *p, int64_t n, int width, Item item) {
    if (n <= 0) return p;
    int T = n < SERIAL_BELOW ? 1 : io_threads();
    if (T == 1) {
        for (int64_t i = 0; i < n; ++i) {
            p = item(p, i);
            *p++ = ' ';
        }
        return p - 1;
    }
    int64_t per = (n + T - 1) / T;
    std::vector<char *> ends(T, nullptr);
    auto chunk = [=, &ends](int k) {
        int64_t lo = std::min<int64_t>(n, k * per), hi = std::min<int64_t>(n, lo + per);
        char *q = p + lo * width;
        for (int64_t i = lo; i < hi; ++i) {
            q = item(q, i);
            *q++ = ' ';
        }
        ends[k] = q;
    };
    std::vector<std::thread> pool;
    pool.reserve(T);
    int started = 0;
    try {
        for (; started < T - 1; ++started) pool.emplace_back(chunk, started);
    } catch (...) {   // no more threads to be had: the caller's thread takes the chunks that are left
    }
    for (int k = started; k < T; ++k) chunk(k);
    for (auto &t : pool) t.join();
    char *w = p;
    for (int k = 0; k < T; ++k) {
        char *src = p + std::min<int64_t>(n, k * per) * width;
        size_t len = size_t(ends[k] - src);
        if (w != src) std::memmove(w, src, len);
        w += len;
    }
    return w - 1;
}

inline int64_t piece_bound(int64_t n_v, int64_t n_t) {
    return 1024 + n_v * int64_t(3 * F64_WIDTH + 4 + F64_WIDTH) + n_t * int64_t(3 * I32_WIDTH + I64_WIDTH + 2);
}

}   // namespace

extern "C" {

int64_t pdn_format_f64(const double *x, int64_t n, char *out, int64_t cap) {
    if (n < 0 || (n > 0 && (!x || !out)) || cap < n * F64_WIDTH) return -PDN_EINVAL;
    return join_items(out, n, F64_WIDTH, [x](char *q, int64_t i) { return put_f64(q, x[i]); }) - out;
}

int64_t pdn_vtu_piece_bound(int64_t n_v, int64_t n_t) {
    return n_v < 0 || n_t < 0 ? -PDN_EINVAL : piece_bound(n_v, n_t);
}

int64_t pdn_vtu_piece(int64_t n_v, const double *xy, const double *voltage, int64_t n_t, const int32_t *tri,
                      char *out, int64_t cap) {
    if (n_v < 0 || n_t < 0 || !out || (n_v > 0 && (!xy || !voltage)) || (n_t > 0 && !tri)) return -PDN_EINVAL;
    if (cap < piece_bound(n_v, n_t)) return -PDN_EINVAL;
    for (int64_t i = 0; i < 3 * n_t; ++i)
        if (tri[i] < 0 || tri[i] >= n_v) return -PDN_EINVAL;
    char *p = out;
    // create_piece, padne/paraview.py:209-231 (lxml pretty_print: two spaces per level, Piece at level 2)
    p = put_bytes(p, "    <Piece NumberOfPoints=\"");
    p = put_i64(p, n_v);
    p = put_bytes(p, "\" NumberOfCells=\"");
    p = put_i64(p, n_t);
    p = put_bytes(p, "\">\n");
    // create_point_data, padne/paraview.py:109-126: potentials in vertex index order
    p = put_bytes(p, "      <PointData Scalars=\"voltage\">\n"
                     "        <DataArray type=\"Float64\" format=\"ascii\" Name=\"voltage\">");
    p = join_items(p, n_v, F64_WIDTH, [voltage](char *q, int64_t i) { return put_f64(q, voltage[i]); });
    p = put_bytes(p, "</DataArray>\n      </PointData>\n");
    // create_points, padne/paraview.py:129-146: (x, -y, 0.0) per vertex
    p = put_bytes(p, "      <Points>\n"
                     "        <DataArray type=\"Float64\" format=\"ascii\" NumberOfComponents=\"3\">");
    p = join_items(p, n_v, 2 * F64_WIDTH + 4, [xy](char *q, int64_t i) {
        q = put_f64(q, xy[2 * i]);
        *q++ = ' ';
        q = put_f64(q, -xy[2 * i + 1]);
        return put_bytes(q, " 0.0");
    });
    p = put_bytes(p, "</DataArray>\n      </Points>\n");
    // create_cells, padne/paraview.py:184-206: connectivity in mesh.faces order, offsets 3(i+1), type 5
    p = put_bytes(p, "      <Cells>\n"
                     "        <DataArray type=\"Int32\" format=\"ascii\" Name=\"connectivity\">");
    p = join_items(p, n_t, 3 * I32_WIDTH, [tri](char *q, int64_t i) {
        q = put_i64(q, tri[3 * i]);
        *q++ = ' ';
        q = put_i64(q, tri[3 * i + 1]);
        *q++ = ' ';
        return put_i64(q, tri[3 * i + 2]);
    });
    p = put_bytes(p, "</DataArray>\n"
                     "        <DataArray type=\"Int32\" format=\"ascii\" Name=\"offsets\">");
    p = join_items(p, n_t, I64_WIDTH, [](char *q, int64_t i) { return put_i64(q, 3 * (i + 1)); });
    p = put_bytes(p, "</DataArray>\n"
                     "        <DataArray type=\"UInt8\" format=\"ascii\" Name=\"types\">");
    if (n_t > 0) {
        for (int64_t i = 0; i < n_t; ++i) {
            *p++ = '5';
            *p++ = ' ';
        }
        --p;
    }
    p = put_bytes(p, "</DataArray>\n      </Cells>\n    </Piece>\n");
    return p - out;
}

}   // extern "C"
