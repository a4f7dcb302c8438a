// Native reader and writer for the reference's dense text format (host code only; lives in the CUDA
// library so that there is one shared object).
//
// The reference stores every dense partition as text, one row per line, space-separated decimal
// floats (util.py:26-30) and reads it back with np.loadtxt (util.py:13-15), which parses a few
// million values per second on one core -- minutes for the partitions this framework is built for.
// eh_load_dense_text maps the file, cuts it into line-aligned chunks, parses the chunks on all
// cores with a hand-written decimal parser (digits accumulated in a 64-bit integer, one scaling by
// a power of ten; strtod only for what it does not cover: more than 19 significant digits, inf /
// nan, hex) and writes float32 or float64 rows with an arbitrary row stride, so the result can be
// the pinned staging buffer of a partition with its padded row stride.
#include <errno.h>
#include <fcntl.h>
#include <math.h>
#include <stdlib.h>
#include <string.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

#include <atomic>
#include <charconv>
#include <chrono>
#include <string>
#include <thread>
#include <vector>

#include "common.cuh"

namespace eh {

static const double kPow10[] = {1e0,  1e1,  1e2,  1e3,  1e4,  1e5,  1e6,  1e7,  1e8,  1e9,  1e10, 1e11,
                                1e12, 1e13, 1e14, 1e15, 1e16, 1e17, 1e18, 1e19, 1e20, 1e21, 1e22};

static const long double kPow10L[] = {1e0L,  1e1L,  1e2L,  1e3L,  1e4L,  1e5L,  1e6L,  1e7L,  1e8L,  1e9L,
                                      1e10L, 1e11L, 1e12L, 1e13L, 1e14L, 1e15L, 1e16L, 1e17L, 1e18L, 1e19L,
                                      1e20L, 1e21L, 1e22L, 1e23L, 1e24L, 1e25L, 1e26L, 1e27L};

// Parses one number at p (p < end, not whitespace).  Returns the position after it, or nullptr on a syntax error.
static inline const char* parse_number(const char* p, const char* end, double* out) {
    const char* start = p;
    bool neg = false;
    if (p < end && (*p == '-' || *p == '+')) { neg = (*p == '-'); ++p; }
    unsigned long long mant = 0;
    int digits = 0, frac = 0;
    bool any = false, exact = true;
    while (p < end && *p >= '0' && *p <= '9') {
        if (digits < 19) { mant = mant * 10 + (unsigned)(*p - '0'); digits += (mant != 0); } else { exact = false; }
        any = true; ++p;
    }
    if (p < end && *p == '.') {
        ++p;
        while (p < end && *p >= '0' && *p <= '9') {
            if (digits < 19) { mant = mant * 10 + (unsigned)(*p - '0'); digits += (mant != 0); ++frac; } else { exact = false; }
            any = true; ++p;
        }
    }
    int ex = 0;
    if (any && p < end && (*p == 'e' || *p == 'E')) {
        const char* q = p + 1;
        bool eneg = false;
        if (q < end && (*q == '-' || *q == '+')) { eneg = (*q == '-'); ++q; }
        if (q < end && *q >= '0' && *q <= '9') {
            int v = 0;
            while (q < end && *q >= '0' && *q <= '9') { if (v < 10000) v = v * 10 + (*q - '0'); ++q; }
            ex = eneg ? -v : v;
            p = q;
        }
    }
    const int e10 = ex - frac;
    // exact fast path: a mantissa below 2^53 scaled by one exactly representable power of ten is correctly rounded
    if (any && exact && mant < (1ull << 53) && e10 >= -22 && e10 <= 22) {
        double v = (double)mant;
        v = e10 < 0 ? v / kPow10[-e10] : v * kPow10[e10];
        *out = neg ? -v : v;
        return p;
    }
    // up to 19 significant digits (the 16-17 digit shortest representations Python writes): one operation in the
    // 64-bit-mantissa extended format, then the rounding to double.  The extended result is within one unit of its
    // last place of the true value, so the second rounding is the correct one unless the 11 bits below double
    // precision sit within one unit of the half-way pattern; those (and anything outside the normal range) go on
    // to the C library.
    if (any && exact && e10 >= -27 && e10 <= 27 && sizeof(long double) >= 10) {
        long double v = (long double)mant;
        v = e10 < 0 ? v / kPow10L[-e10] : v * kPow10L[e10];
        unsigned long long m64;
        memcpy(&m64, &v, 8);
        const unsigned low = (unsigned)(m64 & 0x7ffull);
        const long double av = v < 0 ? -v : v;
        if ((low < 0x3feu || low > 0x402u) && (mant == 0 || (av > 1e-290L && av < 1e290L))) {
            const double dv = (double)v;
            *out = neg ? -dv : dv;
            return p;
        }
    }
    // everything else (long mantissas, big exponents, inf, nan): the C library, on a bounded copy
    char buf[96];
    const char* q = start;
    while (q < end && *q != ' ' && *q != '\t' && *q != '\n' && *q != '\r' && *q != ',') ++q;
    const size_t len = (size_t)(q - start);
    if (len == 0 || len >= sizeof(buf)) return nullptr;
    memcpy(buf, start, len);
    buf[len] = 0;
    char* stop = nullptr;
    const double v = strtod(buf, &stop);
    if (stop != buf + len) return nullptr;
    *out = v;
    return q;
}

struct TextChunk {
    const char* begin;
    const char* end;
    long long first_row;      // filled by the counting pass
    long long rows;
};

// parser threads: all cores (at most 64), or EH_TEXT_THREADS
static unsigned text_threads() {
    const char* e = getenv("EH_TEXT_THREADS");
    if (e && atoi(e) > 0) return (unsigned)std::min(atoi(e), 256);
    return std::max(1u, std::min(std::thread::hardware_concurrency(), 64u));
}

static inline bool is_blank_line(const char* p, const char* e) {
    for (; p < e; ++p)
        if (*p != ' ' && *p != '\t' && *p != '\r') return false;
    return true;
}

}  // namespace eh

// Counts rows and columns of a dense text matrix.  Returns 0 or a negative EH_E* code.
extern "C" int eh_text_shape(const char* path, int64_t* rows, int64_t* cols) {
    using namespace eh;
    EH_REQUIRE(path && rows && cols, "text_shape: null argument");
    const int fd = open(path, O_RDONLY);
    EH_REQUIRE(fd >= 0, "text_shape: cannot open %s: %s", path, strerror(errno));
    struct stat st;
    if (fstat(fd, &st) != 0 || st.st_size == 0) { close(fd); *rows = 0; *cols = 0; return EH_OK; }
    const char* base = (const char*)mmap(nullptr, (size_t)st.st_size, PROT_READ, MAP_PRIVATE | MAP_POPULATE, fd, 0);
    close(fd);
    EH_REQUIRE(base != MAP_FAILED, "text_shape: mmap of %s failed: %s", path, strerror(errno));
    const char* end = base + st.st_size;
    // columns: numbers on the first non-blank line
    const char* p = base;
    long long c = 0;
    while (p < end && c == 0) {
        const char* nl = (const char*)memchr(p, '\n', (size_t)(end - p));
        const char* le = nl ? nl : end;
        const char* q = p;
        while (q < le) {
            while (q < le && (*q == ' ' || *q == '\t' || *q == '\r' || *q == ',')) ++q;
            if (q >= le) break;
            ++c;
            while (q < le && *q != ' ' && *q != '\t' && *q != '\r' && *q != ',') ++q;
        }
        p = nl ? nl + 1 : end;
    }
    // rows: non-blank lines, counted in parallel
    const unsigned nt = text_threads();
    std::vector<long long> cnt(nt, 0);
    std::vector<std::thread> th;
    const size_t sz = (size_t)st.st_size;
    for (unsigned t = 0; t < nt; ++t)
        th.emplace_back([&, t] {
            const char* b = base + sz * t / nt;
            const char* e = base + sz * (t + 1) / nt;
            if (t > 0) {                                    // start after the line that straddles the boundary
                const char* nl = (const char*)memchr(b - 1, '\n', (size_t)(end - (b - 1)));
                b = nl ? nl + 1 : end;
            }
            long long n = 0;
            while (b < e) {
                const char* nl = (const char*)memchr(b, '\n', (size_t)(end - b));
                const char* le = nl ? nl : end;
                if (!is_blank_line(b, le)) ++n;
                b = nl ? nl + 1 : end;
            }
            cnt[t] = n;
        });
    for (auto& x : th) x.join();
    long long r = 0;
    for (auto v : cnt) r += v;
    munmap((void*)base, sz);
    *rows = r;
    *cols = c;
    return EH_OK;
}

// Parses a rows x cols text matrix into out (float32 if elem_bytes == 4, float64 if 8) with row stride ld
// elements (ld >= cols; the padding is left untouched).  Returns 0, or EH_EINVAL with a message that names
// the offending line.
extern "C" int eh_load_dense_text(const char* path, int64_t rows, int64_t cols, void* out, int64_t ld, int elem_bytes) {
    using namespace eh;
    EH_REQUIRE(path && out && rows >= 0 && cols > 0 && ld >= cols, "load_dense_text: bad argument");
    EH_REQUIRE(elem_bytes == 4 || elem_bytes == 8, "load_dense_text: elem_bytes must be 4 or 8");
    if (rows == 0) return EH_OK;
    const int fd = open(path, O_RDONLY);
    EH_REQUIRE(fd >= 0, "load_dense_text: cannot open %s: %s", path, strerror(errno));
    struct stat st;
    if (fstat(fd, &st) != 0 || st.st_size == 0) { close(fd); set_error("load_dense_text: %s is empty", path); return EH_EINVAL; }
    const size_t sz = (size_t)st.st_size;
    const char* base = (const char*)mmap(nullptr, sz, PROT_READ, MAP_PRIVATE | MAP_POPULATE, fd, 0);
    close(fd);
    EH_REQUIRE(base != MAP_FAILED, "load_dense_text: mmap of %s failed: %s", path, strerror(errno));
    const char* end = base + sz;
    const unsigned nt = text_threads();
    std::vector<TextChunk> ch(nt);
    for (unsigned t = 0; t < nt; ++t) {
        const char* b = base + sz * t / nt;
        if (t > 0) {
            const char* nl = (const char*)memchr(b - 1, '\n', (size_t)(end - (b - 1)));
            b = nl ? nl + 1 : end;
        }
        ch[t].begin = b;
        if (t > 0) ch[t - 1].end = b;
    }
    ch[nt - 1].end = end;
    const bool dbg = getenv("EH_TEXT_DEBUG") != nullptr;
    auto now = [] { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); };
    const double t_a = now();
    // pass 1: rows per chunk -> first row of every chunk
    {
        std::vector<std::thread> th;
        for (unsigned t = 0; t < nt; ++t)
            th.emplace_back([&, t] {
                long long n = 0;
                const char* b = ch[t].begin;
                while (b < ch[t].end) {
                    const char* nl = (const char*)memchr(b, '\n', (size_t)(ch[t].end - b));
                    const char* le = nl ? nl : ch[t].end;
                    if (!is_blank_line(b, le)) ++n;
                    b = nl ? nl + 1 : ch[t].end;
                }
                ch[t].rows = n;
            });
        for (auto& x : th) x.join();
    }
    long long total = 0;
    for (unsigned t = 0; t < nt; ++t) { ch[t].first_row = total; total += ch[t].rows; }
    if (dbg)
        for (unsigned t = 0; t < nt; ++t)
            fprintf(stderr, "[textio] chunk %u: %zu bytes, %lld rows\n", t, (size_t)(ch[t].end - ch[t].begin), ch[t].rows);
    if (total != rows) {
        munmap((void*)base, sz);
        set_error("load_dense_text: %s has %lld rows, expected %lld", path, total, (long long)rows);
        return EH_EINVAL;
    }
    const double t_b = now();
    // pass 2: parse
    std::atomic<long long> bad_row{-1};
    {
        std::vector<std::thread> th;
        for (unsigned t = 0; t < nt; ++t)
            th.emplace_back([&, t] {
                long long r = ch[t].first_row;
                const char* b = ch[t].begin;
                while (b < ch[t].end && bad_row.load(std::memory_order_relaxed) < 0) {
                    const char* nl = (const char*)memchr(b, '\n', (size_t)(ch[t].end - b));
                    const char* le = nl ? nl : ch[t].end;
                    if (!is_blank_line(b, le)) {
                        const char* q = b;
                        long long c = 0;
                        while (true) {
                            while (q < le && (*q == ' ' || *q == '\t' || *q == '\r' || *q == ',')) ++q;
                            if (q >= le) break;
                            double v;
                            const char* nx = (c < cols) ? parse_number(q, le, &v) : nullptr;
                            if (!nx) { c = -1; break; }
                            if (elem_bytes == 4) ((float*)out)[r * ld + c] = (float)v;
                            else ((double*)out)[r * ld + c] = v;
                            ++c;
                            q = nx;
                        }
                        if (c != cols) { long long exp = -1; bad_row.compare_exchange_strong(exp, r); return; }
                        ++r;
                    }
                    b = nl ? nl + 1 : ch[t].end;
                }
            });
        for (auto& x : th) x.join();
    }
    if (dbg) fprintf(stderr, "[textio] %u threads: count %.3f s, parse %.3f s\n", nt, t_b - t_a, now() - t_b);
    munmap((void*)base, sz);
    if (bad_row.load() >= 0) {
        set_error("load_dense_text: %s: row %lld does not hold exactly %lld numbers", path, bad_row.load() + 1, (long long)cols);
        return EH_EINVAL;
    }
    return EH_OK;
}

// ------------------------------------------------------------------------------------------
// Writer.  The reference writes a partition with  " ".join(str(x) for x in row)  (util.py:26-30): every value
// is Python's repr of a double -- the shortest decimal string that reads back to the same double, in positional
// notation with at least one digit after the point for 1e-4 <= |x| < 1e16 and as d[.ddd]e+XX otherwise.  A
// million values per second in Python; here std::to_chars supplies the shortest digits and format_repr lays them
// out by the same rule, rows are formatted by all cores in slabs and written in order, so the file is
// byte-identical to the reference's.
// ------------------------------------------------------------------------------------------
namespace eh {

// Appends repr(v) at p (at least 32 bytes free); returns the new end.
static inline char* format_repr(char* p, double v) {
    if (std::isnan(v)) { memcpy(p, "nan", 3); return p + 3; }
    if (std::isinf(v)) { if (v < 0) *p++ = '-'; memcpy(p, "inf", 3); return p + 3; }
    char tmp[40];
    const auto r = std::to_chars(tmp, tmp + sizeof(tmp), v, std::chars_format::scientific);   // [-]d[.ddd]e[+-]XX
    const char* q = tmp;
    if (*q == '-') { *p++ = '-'; ++q; }
    char digits[24];
    int nd = 0;
    digits[nd++] = *q++;
    if (*q == '.') { ++q; while (*q != 'e') digits[nd++] = *q++; }
    ++q;                                                     // 'e'
    const bool eneg = (*q == '-');
    ++q;
    int ex = 0;
    while (q < r.ptr) ex = ex * 10 + (*q++ - '0');
    if (eneg) ex = -ex;
    if (ex < -4 || ex >= 16) {                               // float_repr_style 'short', format code 'r'
        *p++ = digits[0];
        if (nd > 1) { *p++ = '.'; memcpy(p, digits + 1, (size_t)(nd - 1)); p += nd - 1; }
        *p++ = 'e';
        *p++ = ex < 0 ? '-' : '+';
        int a = ex < 0 ? -ex : ex;
        if (a >= 100) { *p++ = (char)('0' + a / 100); a %= 100; }
        *p++ = (char)('0' + a / 10);
        *p++ = (char)('0' + a % 10);
        return p;
    }
    if (ex < 0) {                                            // 0.000ddd
        *p++ = '0'; *p++ = '.';
        for (int i = 0; i < -ex - 1; ++i) *p++ = '0';
        memcpy(p, digits, (size_t)nd);
        return p + nd;
    }
    const int ip = ex + 1;                                   // digits before the point
    if (nd <= ip) {
        memcpy(p, digits, (size_t)nd); p += nd;
        for (int i = nd; i < ip; ++i) *p++ = '0';
        *p++ = '.'; *p++ = '0';
        return p;
    }
    memcpy(p, digits, (size_t)ip); p += ip;
    *p++ = '.';
    memcpy(p, digits + ip, (size_t)(nd - ip));
    return p + (nd - ip);
}

}  // namespace eh

extern "C" int eh_format_repr(double v, char* out32) {
    if (!out32) return EH_EINVAL;
    char* e = eh::format_repr(out32, v);
    *e = 0;
    return (int)(e - out32);
}

extern "C" int eh_save_dense_text(const char* path, const double* data, int64_t rows, int64_t cols, int64_t ld) {
    using namespace eh;
    EH_REQUIRE(path && (data || rows == 0) && rows >= 0 && cols > 0 && ld >= cols, "save_dense_text: bad argument");
    FILE* f = fopen(path, "wb");
    EH_REQUIRE(f != nullptr, "save_dense_text: cannot create %s: %s", path, strerror(errno));
    const unsigned nt = text_threads();
    // slabs of rows: about 4 MB of text per thread and wave
    long long slab = (long long)((4u << 20) / (size_t)(cols * 20 + 1));
    if (slab < 1) slab = 1;
    std::vector<std::string> buf(nt);
    bool ok = true;
    for (long long r0 = 0; r0 < rows && ok; r0 += slab * nt) {
        std::vector<std::thread> th;
        for (unsigned t = 0; t < nt; ++t) {
            const long long a = r0 + (long long)t * slab, b = std::min<long long>(a + slab, rows);
            buf[t].clear();
            if (a >= b) continue;
            th.emplace_back([&, t, a, b] {
                std::string& s = buf[t];
                s.resize((size_t)(b - a) * (size_t)(cols * 26 + 1));
                char* p = &s[0];
                for (long long r = a; r < b; ++r) {
                    const double* row = data + r * ld;
                    for (long long c = 0; c < cols; ++c) {
                        if (c) *p++ = ' ';
                        p = format_repr(p, row[c]);
                    }
                    *p++ = '\n';
                }
                s.resize((size_t)(p - &s[0]));
            });
        }
        for (auto& x : th) x.join();
        for (unsigned t = 0; t < nt && ok; ++t)
            if (!buf[t].empty()) ok = fwrite(buf[t].data(), 1, buf[t].size(), f) == buf[t].size();
    }
    const bool closed = fclose(f) == 0;
    if (!ok || !closed) { set_error("save_dense_text: write to %s failed: %s", path, strerror(errno)); return EH_EINVAL; }
    return EH_OK;
}
