// scene.cpp — OBJ/MTL loading with the reference's semantics (scene.cpp:26-109) and the GPU hand-off.
//
// Loader rules kept from the reference: whitespace-separated token stream; `mtllib` is resolved against
// Config::model_path and read as newmtl blocks of Kd / Ke / Ks / Ns (everything else ignored), one material per
// newmtl in file order; `v` x y z; `vt` skipped; `vn` x y z; `f a/b/c a/b/c a/b/c` triangles only, 1-based; when
// the first corner carries a normal index the plane is re-derived from that `vn`; `usemtl <int>` selects a
// material by its ordinal.  An unopenable OBJ is fatal (exit(1)), as scene.cpp:32-35.
#include "scene.h"

#include <fcntl.h>
#include <sys/stat.h>
#include <unistd.h>

#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <algorithm>
#include <chrono>
#include <iostream>
#include <memory>
#include <thread>
#include <utility>

using ptm::vec3;
using ptm::vec4;

Scene::Scene(std::vector<std::vector<vec3>> &color_map, std::vector<std::vector<vec3>> &color2_map,
             std::vector<std::vector<int>> &samples_count)
    : color_map_(color_map), color2_map_(color2_map), samples_count_(samples_count) {
    if (!Config::get().skybox.empty()) skybox_.load(Config::get().skybox);
}

int Scene::AddMaterial(const MaterialCharacteristics &ch) {
    materials_.push_back(Factory(ch));
    return static_cast<int>(materials_.size()) - 1;
}

void Scene::AddTriangle(Triangle triangle) { triangles_.push_back(triangle); }

namespace {
// "a/b/c" -> three 0-based indices, -1 where a field is missing or empty (atoi("") - 1, scene.cpp:91-96)
void corner(const char *b, const char *e, int out[3]) {
    for (int k = 0; k < 3; ++k) {
        if (b > e) { out[k] = -1; continue; }  // no more fields
        // atoi: optional sign, digits; anything else ends the number (the token's '/' or its end)
        const char *q = b;
        bool neg = false;
        if (q < e && (*q == '-' || *q == '+')) neg = *q++ == '-';
        long v = 0;
        while (q < e && *q >= '0' && *q <= '9') v = v * 10 + (*q++ - '0');
        out[k] = static_cast<int>(neg ? -v : v) - 1;
        while (q < e && *q != '/') ++q;  // rest of the field
        b = q + 1;                        // past the '/', or past the end: later fields are missing
    }
}

void read_mtl(const std::string &path, Scene &scene) {
    std::ifstream in(path);
    std::string tok = "1";
    // the reference emits one material per pass of its outer loop, including a default-valued one when the file
    // is missing or ends right after the last block's values (scene.cpp:45-70)
    while (!in.eof()) {
        MaterialCharacteristics ch;
        while (!in.eof() && tok != "newmtl") in >> tok;
        in >> tok;  // the material's name (unused: usemtl is an ordinal)
        while (!in.eof() && tok != "newmtl") {
            if (tok == "Kd") in >> ch.Kd.x >> ch.Kd.y >> ch.Kd.z;
            else if (tok == "Ke") in >> ch.Ke.x >> ch.Ke.y >> ch.Ke.z;
            else if (tok == "Ks") in >> ch.Ks.x >> ch.Ks.y >> ch.Ks.z;
            else if (tok == "Ns") in >> ch.Ns;
            else if (tok == "Ni") in >> ch.Ni;                                  // extension (the reference skips these)
            else if (tok == "Tr") in >> ch.Tr;
            else if (tok == "d") { float d = 1; in >> d; ch.Tr = 1 - d; }
            in >> tok;
        }
        scene.AddMaterial(ch);
    }
}
}  // namespace

namespace {
// The OBJ as one buffer, read token by token.  Same token stream as the reference's `file >> input` (whitespace
// separated; scene.cpp:37-38), same numeric conversion (strtof / strtol are what libstdc++'s operator>> ends in), but
// no locale, sentry and stream-state work per token: a 1 M-triangle OBJ (40 MB) loads in a fraction of a second
// instead of the seconds the stream extraction takes.
// strtof for the common case, bit for bit.  A decimal with at most 15 significant digits and a power of ten within
// 10^+-22 converts to double with ONE rounding (digits and power are exact doubles: Clinger's fast path), and that
// double rounds to the float strtof returns unless it sits exactly on the midpoint of two floats — if the exact value
// and its double were on different sides of a midpoint, the midpoint, itself a double, would be the nearer double.
// On a midpoint bit pattern, or for anything else this does not take (hex, inf / nan, long digit strings, large
// exponents), the caller calls strtof.  p[...] is readable up to a terminating non-numeric character.
inline bool fast_strtof(const char *p, float *out, const char **endp) {
    static const double kPow10[23] = {1e0,  1e1,  1e2,  1e3,  1e4,  1e5,  1e6,  1e7,  1e8,  1e9,  1e10, 1e11,
                                      1e12, 1e13, 1e14, 1e15, 1e16, 1e17, 1e18, 1e19, 1e20, 1e21, 1e22};
    const char *q = p;
    bool neg = false;
    if (*q == '-' || *q == '+') neg = *q++ == '-';
    if (!((*q >= '0' && *q <= '9') || *q == '.')) return false;
    if (*q == '0' && (q[1] == 'x' || q[1] == 'X')) return false;
    uint64_t mant = 0;
    int sig = 0, frac = 0;
    bool any = false;
    for (; *q >= '0' && *q <= '9'; ++q) {
        any = true;
        if (sig || *q != '0') {
            if (sig == 15) return false;
            mant = mant * 10 + static_cast<uint64_t>(*q - '0');
            ++sig;
        }
    }
    if (*q == '.') {
        for (++q; *q >= '0' && *q <= '9'; ++q) {
            any = true;
            ++frac;
            if (sig || *q != '0') {
                if (sig == 15) return false;
                mant = mant * 10 + static_cast<uint64_t>(*q - '0');
                ++sig;
            }
        }
    }
    if (!any) return false;
    int e10 = 0;
    if (*q == 'e' || *q == 'E') {
        const char *r = q + 1;
        bool eneg = false;
        if (*r == '-' || *r == '+') eneg = *r++ == '-';
        if (*r >= '0' && *r <= '9') {
            int digits = 0;
            for (; *r >= '0' && *r <= '9'; ++r) {
                if (++digits > 4) return false;
                e10 = e10 * 10 + (*r - '0');
            }
            if (eneg) e10 = -e10;
            q = r;
        }  // no digits: the number ends before the 'e', as for strtof
    }
    float f;
    if (mant == 0) {
        f = 0.0f;
    } else {
        const int k = e10 - frac;
        if (k < -22 || k > 22) return false;
        const double d = k < 0 ? static_cast<double>(mant) / kPow10[-k] : static_cast<double>(mant) * kPow10[k];
        uint64_t bits;
        std::memcpy(&bits, &d, sizeof bits);
        if ((bits & 0x1fffffffull) == 0x10000000ull) return false;  // a float midpoint: let strtof look at the digits
        f = static_cast<float>(d);
    }
    *out = neg ? -f : f;
    *endp = q;
    return true;
}

class TokenStream {
    const char *p_, *end_;  // the buffer is followed by one ' ': strtof / strtol stop there

public:
    TokenStream(const char *begin, const char *end) : p_(begin), end_(end) {}
    static bool space(char c) { return c == ' ' || c == '\n' || c == '\t' || c == '\r' || c == '\v' || c == '\f'; }
    // start of the next token (end of the buffer if there is none)
    const char *peek() {
        while (p_ < end_ && space(*p_)) ++p_;
        return p_;
    }
    // next token as [b, e); false at end of file
    bool next(const char *&b, const char *&e) {
        while (p_ < end_ && space(*p_)) ++p_;
        if (p_ >= end_) return false;
        b = p_;
        while (p_ < end_ && !space(*p_)) ++p_;
        e = p_;
        return true;
    }
    float next_float() {  // like `file >> f`: skips whitespace, converts the longest numeric prefix
        while (p_ < end_ && space(*p_)) ++p_;
        float fast;
        const char *fe;
        if (fast_strtof(p_, &fast, &fe)) {
            p_ = fe;
            return fast;
        }
        char *q = nullptr;
        const float f = std::strtof(p_, &q);
        if (q == p_) {  // not a number: the stream would fail here; consume the token, value 0
            const char *b, *e;
            next(b, e);
            return 0.0f;
        }
        p_ = q;
        return f;
    }
    int next_int() {
        while (p_ < end_ && space(*p_)) ++p_;
        char *q = nullptr;
        const long v = std::strtol(p_, &q, 10);
        if (q == p_) { const char *b, *e; next(b, e); return 0; }
        p_ = q;
        return static_cast<int>(v);
    }
};
// The whole file followed by one ' ' (strtof / strtol stop there), read with pread by a few threads into memory that
// is not zero-filled first (a 135 MB model: 35 ms instead of 220 ms through ifstream into a resized string).
struct FileText {
    std::unique_ptr<char[]> data;
    size_t size = 0;  // without the trailing ' '
    bool ok = false;
};
FileText slurp(const std::string &path) {
    FileText t;
    const int fd = ::open(path.c_str(), O_RDONLY);
    if (fd < 0) return t;
    struct stat st;
    if (::fstat(fd, &st) != 0 || st.st_size < 0) { ::close(fd); return t; }
    t.size = static_cast<size_t>(st.st_size);
    t.data.reset(new char[t.size + 1]);
    const size_t parts = t.size >= (size_t{8} << 20) ? std::min<size_t>(std::max(1u, std::thread::hardware_concurrency()), 8) : 1;
    std::vector<char> good(parts, 1);
    auto read_part = [&](size_t i) {
        size_t off = t.size / parts * i;
        const size_t stop = i + 1 == parts ? t.size : t.size / parts * (i + 1);
        while (off < stop) {
            const ssize_t n = ::pread(fd, t.data.get() + off, stop - off, static_cast<off_t>(off));
            if (n <= 0) { good[i] = 0; return; }
            off += static_cast<size_t>(n);
        }
    };
    std::vector<std::thread> th;
    for (size_t i = 1; i < parts; ++i) th.emplace_back(read_part, i);
    read_part(0);
    for (std::thread &x : th) x.join();
    ::close(fd);
    t.data[t.size] = ' ';
    t.ok = std::all_of(good.begin(), good.end(), [](char g) { return g != 0; });
    return t;
}
inline bool is_tok(const char *b, const char *e, const char *lit) {
    const size_t n = std::strlen(lit);
    return static_cast<size_t>(e - b) == n && std::memcmp(b, lit, n) == 0;
}
}  // namespace

void pth_check_fast_strtof_impl(const char *text, int64_t *counts) {
    counts[0] = counts[1] = counts[2] = 0;
    const char *p = text;
    while (true) {
        while (*p && TokenStream::space(*p)) ++p;
        if (!*p) break;
        ++counts[0];
        float fast = 0;
        const char *fe = nullptr;
        if (fast_strtof(p, &fast, &fe)) {
            ++counts[1];
            char *q = nullptr;
            const float f = std::strtof(p, &q);
            if (q != fe || std::memcmp(&f, &fast, sizeof f) != 0) ++counts[2];
        }
        while (*p && !TokenStream::space(*p)) ++p;
    }
}

// ---- binary scene cache (extension; -CACHE <dir>) ------------------------------------------------------------
// What LoadModel produces — triangles as constructed (vertices, plane, area, material ordinal) and the material
// table — written as flat arrays next to a validity record: size and modification time of the OBJ and of every MTL
// it pulled in.  A stale, truncated or foreign file is ignored and rewritten; a cache that cannot be written is not
// an error.  The 1 000 014-triangle model: 60 MB of cache read in tens of milliseconds instead of parsing 70 MB of
// text.
namespace {
struct FileStamp {
    int64_t size = -1, mtime_ns = 0;
    bool operator==(const FileStamp &o) const { return size == o.size && mtime_ns == o.mtime_ns; }
};
FileStamp stamp_of(const std::string &path) {
    struct stat st;
    FileStamp s;
    if (::stat(path.c_str(), &st) == 0) {
        s.size = static_cast<int64_t>(st.st_size);
        s.mtime_ns = static_cast<int64_t>(st.st_mtim.tv_sec) * 1000000000ll + st.st_mtim.tv_nsec;
    }
    return s;
}
const char kCacheMagic[8] = {'P', 'T', 'B', '2', 'S', 'C', '0', '1'};
struct CacheHeader {
    char magic[8];
    int64_t n_triangles, n_materials, n_mtl_files;
    FileStamp obj;
};
template <class T> bool put(std::FILE *f, const T *p, size_t n) { return std::fwrite(p, sizeof(T), n, f) == n; }
template <class T> bool get(std::FILE *f, T *p, size_t n) { return std::fread(p, sizeof(T), n, f) == n; }
std::string cache_name(const std::string &dir, const std::string &obj_path) {
    const size_t slash = obj_path.find_last_of('/');
    std::string d = dir;
    if (!d.empty() && d.back() != '/') d += '/';
    return d + (slash == std::string::npos ? obj_path : obj_path.substr(slash + 1)) + ".ptscene";
}
}  // namespace

void Scene::SaveCache(const std::string &cache_file, const std::string &obj_path) const {
    const std::string tmp = cache_file + ".tmp";
    std::FILE *f = std::fopen(tmp.c_str(), "wb");
    if (!f) return;
    CacheHeader h;
    std::memcpy(h.magic, kCacheMagic, 8);
    h.n_triangles = static_cast<int64_t>(triangles_.size());
    h.n_materials = static_cast<int64_t>(materials_.size());
    h.n_mtl_files = static_cast<int64_t>(mtl_files_.size());
    h.obj = stamp_of(obj_path);
    bool ok = put(f, &h, 1);
    for (const std::string &m : mtl_files_) {
        const int64_t len = static_cast<int64_t>(m.size());
        const FileStamp st = stamp_of(m);
        ok = ok && put(f, &len, 1) && put(f, m.data(), m.size()) && put(f, &st, 1);
    }
    const PtScene s = abi();  // the flat arrays
    const size_t n = triangles_.size(), m = materials_.size();
    ok = ok && put(f, s.vertices, 9 * n) && put(f, s.planes, 4 * n) && put(f, s.squares, n) && put(f, s.material_ids, n) &&
         put(f, s.materials, 10 * m) && put(f, mats_ext_.data(), 2 * m);
    ok = (std::fclose(f) == 0) && ok;
    if (!ok || std::rename(tmp.c_str(), cache_file.c_str()) != 0) std::remove(tmp.c_str());
}

bool Scene::LoadCache(const std::string &cache_file, const std::string &obj_path) {
    std::FILE *f = std::fopen(cache_file.c_str(), "rb");
    if (!f) return false;
    CacheHeader h;
    bool ok = get(f, &h, 1) && std::memcmp(h.magic, kCacheMagic, 8) == 0 && h.obj == stamp_of(obj_path) &&
              h.n_triangles >= 0 && h.n_triangles <= 0x7fffffff && h.n_materials >= 0 && h.n_materials < (1 << 24) &&
              h.n_mtl_files >= 0 && h.n_mtl_files < 4096;
    std::vector<std::string> mtls;
    for (int64_t k = 0; ok && k < h.n_mtl_files; ++k) {
        int64_t len = 0;
        FileStamp st;
        ok = get(f, &len, 1) && len >= 0 && len < 65536;
        std::string name(ok ? static_cast<size_t>(len) : 0, '\0');
        ok = ok && get(f, &name[0], name.size()) && get(f, &st, 1) && st == stamp_of(name);
        mtls.push_back(name);
    }
    if (!ok) { std::fclose(f); return false; }
    const size_t n = static_cast<size_t>(h.n_triangles), m = static_cast<size_t>(h.n_materials);
    {   // the payload the header promises must be exactly what is left of the file: checked BEFORE anything is sized
        // from the header (a damaged cache with the right magic and stamp is ignored, not a bad_alloc)
        const long at = std::ftell(f);
        const bool seek_ok = at >= 0 && std::fseek(f, 0, SEEK_END) == 0;
        const long end = seek_ok ? std::ftell(f) : -1;
        const unsigned long long payload = 4ull * (9 * n + 4 * n + n + n + 10 * m + 2 * m);
        if (!seek_ok || end < at || static_cast<unsigned long long>(end - at) != payload || std::fseek(f, at, SEEK_SET) != 0) {
            std::fclose(f);
            return false;
        }
    }
    std::vector<float> v(9 * n), p(4 * n), sq(n), mt(10 * m), ext(2 * m);
    std::vector<int32_t> ids(n);
    ok = get(f, v.data(), v.size()) && get(f, p.data(), p.size()) && get(f, sq.data(), n) && get(f, ids.data(), n) &&
         get(f, mt.data(), mt.size()) && get(f, ext.data(), ext.size());
    char extra;
    ok = ok && std::fread(&extra, 1, 1, f) == 0;  // nothing may follow
    std::fclose(f);
    if (!ok) return false;
    for (size_t j = 0; j < m; ++j) {
        MaterialCharacteristics ch;
        const float *r = &mt[10 * j];
        ch.Ke = vec3(r[0], r[1], r[2]); ch.Kd = vec3(r[3], r[4], r[5]); ch.Ks = vec3(r[6], r[7], r[8]);
        ch.Ns = r[9]; ch.Ni = ext[2 * j]; ch.Tr = ext[2 * j + 1];
        AddMaterial(ch);
    }
    triangles_.reserve(triangles_.size() + n);
    for (size_t i = 0; i < n; ++i) triangles_.emplace_back(&v[9 * i], &p[4 * i], sq[i], ids[i]);
    mtl_files_ = mtls;
    return true;
}

// ---- parallel parse ----------------------------------------------------------------------------
// The reference's loader is a state machine over whitespace-separated tokens: a keyword token, then its arguments,
// wherever the line breaks are; every other token is skipped.  A large file is cut after line breaks into one piece
// per thread, and every piece is parsed from its first token as if a statement began there, running past its end to
// finish a statement it has started.  That guess is checked, not trusted: the pieces are consistent when each one
// stopped exactly at the token the next one started from; then the sequential machine would have gone through the
// same states.  Vertex and normal indices are resolved after a prefix sum over the pieces, with the bound the
// sequential loader's `positions.at()` would have applied at that point of the file (no forward references).
// Anything else — an inconsistent cut, an index out of range, a face cut short by the end of the file — falls back to
// the sequential loop, which then behaves (and throws) exactly as before.
namespace {
struct FaceRec {
    int idx[3][3];
    uint32_t npos, nnrm;  // vertices / normals of THIS piece defined before the face
};
struct Piece {
    const char *begin = nullptr, *limit = nullptr;     // statements starting in [begin, limit) are this piece's
    const char *first_tok = nullptr, *end_tok = nullptr;
    std::vector<vec4> positions;
    std::vector<vec3> normals;
    std::vector<FaceRec> faces;
    std::vector<std::pair<size_t, int>> usemtl;         // (faces before the statement, ordinal)
    std::vector<std::string> mtllibs;
    bool truncated = false;                             // the file ended inside a statement
};

void parse_piece(Piece &pc, const char *file_end) {
    // room for the densest files (short `v` / `f` lines) up front: untouched pages cost nothing, regrowth does
    const size_t bytes = static_cast<size_t>(pc.limit - pc.begin);
    pc.positions.reserve(bytes / 12);
    pc.faces.reserve(bytes / 10);
    TokenStream ts(pc.begin, file_end);
    pc.first_tok = ts.peek();
    const char *b, *e;
    while (ts.peek() < pc.limit && ts.next(b, e)) {
        if (is_tok(b, e, "mtllib")) {
            if (!ts.next(b, e)) { pc.truncated = true; break; }
            pc.mtllibs.emplace_back(b, e);
        } else if (is_tok(b, e, "v")) {
            vec4 p(1);
            p.x = ts.next_float(); p.y = ts.next_float(); p.z = ts.next_float();
            pc.positions.push_back(p);
        } else if (is_tok(b, e, "vt")) {
            ts.next_float(); ts.next_float();
        } else if (is_tok(b, e, "vn")) {
            vec3 n;
            n.x = ts.next_float(); n.y = ts.next_float(); n.z = ts.next_float();
            pc.normals.push_back(n);
        } else if (is_tok(b, e, "f")) {
            FaceRec f;
            for (int i = 0; i < 3; ++i) {
                if (!ts.next(b, e)) { pc.truncated = true; b = e = ""; }
                corner(b, e, f.idx[i]);
            }
            f.npos = static_cast<uint32_t>(pc.positions.size());
            f.nnrm = static_cast<uint32_t>(pc.normals.size());
            pc.faces.push_back(f);
        } else if (is_tok(b, e, "usemtl")) {
            pc.usemtl.emplace_back(pc.faces.size(), ts.next_int());
        }
    }
    pc.end_tok = ts.peek();
}

struct Lap {  // PT_PROFILE=1: where LoadModel spends its time, on stderr
    bool on = std::getenv("PT_PROFILE") != nullptr;
    std::chrono::steady_clock::time_point t = std::chrono::steady_clock::now();
    void operator()(const char *what) {
        if (!on) return;
        const auto n = std::chrono::steady_clock::now();
        std::fprintf(stderr, "[LoadModel] %-18s %8.2f ms\n", what, std::chrono::duration<double, std::milli>(n - t).count());
        t = n;
    }
};

template <class F> void run_threads(size_t n, F f) {
    std::vector<std::thread> th;
    for (size_t i = 1; i < n; ++i) th.emplace_back([&f, i] { f(i); });
    f(0);
    for (std::thread &t : th) t.join();
}
}  // namespace

bool Scene::LoadModelParallel(const char *text, size_t size) {
    // tuning knobs for the tests: PT_LOADER_THREADS (0 / 1 = sequential), PT_LOADER_MIN_BYTES (smallest file cut up)
    const char *env_t = std::getenv("PT_LOADER_THREADS"), *env_b = std::getenv("PT_LOADER_MIN_BYTES");
    size_t threads = env_t ? static_cast<size_t>(std::atol(env_t)) : std::min<size_t>(std::thread::hardware_concurrency(), 32);
    const size_t min_bytes = env_b ? static_cast<size_t>(std::atol(env_b)) : (size_t{4} << 20);
    if (threads < 2 || size < min_bytes || size < 2 * threads) return false;
    const char *base = text, *end = base + size;
    std::vector<Piece> pieces(threads);
    const char *cut = base;
    for (size_t i = 0; i < threads; ++i) {
        pieces[i].begin = cut;
        const char *want = base + size / threads * (i + 1);
        if (i + 1 == threads || want >= end) {
            cut = end;
        } else {
            if (want < cut) want = cut;
            const void *nl = std::memchr(want, '\n', static_cast<size_t>(end - want));
            cut = nl ? static_cast<const char *>(nl) + 1 : end;
        }
        pieces[i].limit = cut;
    }
    Lap lap;
    run_threads(threads, [&](size_t i) { parse_piece(pieces[i], end); });
    lap("parse pieces");
    // the guess behind the cuts: every piece stopped at the token the next one started from
    for (size_t i = 0; i < threads; ++i) {
        if (pieces[i].truncated) return false;
        if (i + 1 < threads && pieces[i].end_tok != pieces[i + 1].first_tok) return false;
    }
    // prefix sums, then the bounds `positions.at()` / `normals.at()` would have checked at that point of the file
    std::vector<size_t> pos0(threads + 1, 0), nrm0(threads + 1, 0), tri0(threads + 1, 0);
    for (size_t i = 0; i < threads; ++i) {
        pos0[i + 1] = pos0[i] + pieces[i].positions.size();
        nrm0[i + 1] = nrm0[i] + pieces[i].normals.size();
        tri0[i + 1] = tri0[i] + pieces[i].faces.size();
    }
    std::vector<char> bad(threads, 0);
    run_threads(threads, [&](size_t i) {
        for (const FaceRec &f : pieces[i].faces) {
            const long long np = static_cast<long long>(pos0[i] + f.npos), nn = static_cast<long long>(nrm0[i] + f.nnrm);
            for (int k = 0; k < 3; ++k)
                if (f.idx[k][0] < 0 || f.idx[k][0] >= np) bad[i] = 1;
            if (f.idx[0][2] >= nn) bad[i] = 1;
        }
    });
    for (char b : bad)
        if (b) return false;
    lap("check");
    // from here on nothing can fail: materials in file order, then the triangles
    for (const Piece &pc : pieces)
        for (const std::string &name : pc.mtllibs) {
            mtl_files_.push_back(Config::get().model_path + name);
            read_mtl(mtl_files_.back(), *this);
        }
    lap("materials");
    std::vector<int> mat_in(threads + 1, 0);  // current_material at the start of each piece
    for (size_t i = 0; i < threads; ++i) mat_in[i + 1] = pieces[i].usemtl.empty() ? mat_in[i] : pieces[i].usemtl.back().second;
    // global (file order) vertex / normal index -> the piece that parsed it
    auto position = [&](int g) -> const vec4 & {
        const size_t p = static_cast<size_t>(std::upper_bound(pos0.begin() + 1, pos0.end(), static_cast<size_t>(g)) - pos0.begin() - 1);
        return pieces[p].positions[static_cast<size_t>(g) - pos0[p]];
    };
    auto normal = [&](int g) -> const vec3 & {
        const size_t p = static_cast<size_t>(std::upper_bound(nrm0.begin() + 1, nrm0.end(), static_cast<size_t>(g)) - nrm0.begin() - 1);
        return pieces[p].normals[static_cast<size_t>(g) - nrm0[p]];
    };
    // room first (one serial fill: std::vector cannot grow without constructing), then every piece writes its own range
    const size_t t_base = triangles_.size();
    const float zero9[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0}, zero4[4] = {0, 0, 0, 0};
    triangles_.resize(t_base + tri0[threads], Triangle(zero9, zero4, 0.0f, 0));
    lap("room");
    run_threads(threads, [&](size_t i) {
        const Piece &pc = pieces[i];
        Triangle *dst = triangles_.data() + t_base + tri0[i];
        int material = mat_in[i];
        size_t ev = 0;
        for (size_t k = 0; k < pc.faces.size(); ++k) {
            while (ev < pc.usemtl.size() && pc.usemtl[ev].first <= k) material = pc.usemtl[ev++].second;
            const FaceRec &f = pc.faces[k];
            Triangle t({position(f.idx[0][0]), position(f.idx[1][0]), position(f.idx[2][0])}, material);
            if (f.idx[0][2] >= 0) t.SetNormal(normal(f.idx[0][2]));
            dst[k] = t;
        }
    });
    lap("triangles");
    parsed_in_pieces = static_cast<int>(threads);
    return true;
}

void Scene::LoadModel(std::string path) {
    std::ifstream file(path, std::ios::binary);
    if (!file.is_open()) {
        std::cerr << "Impossible to open the file !" << std::endl;
        std::exit(1);
    }
    loaded_from_cache = false;
    parsed_in_pieces = 0;
    const std::string &cache_dir = Config::get().cache;
    const bool fresh = triangles_.empty() && materials_.empty();  // the image describes a whole scene, not an increment
    if (!cache_dir.empty() && fresh && LoadCache(cache_name(cache_dir, path), path)) {
        loaded_from_cache = true;
        return;
    }
    mtl_files_.clear();
    Lap lap;
    file.close();
    const FileText text = slurp(path);
    if (!text.ok) {
        std::cerr << "Impossible to open the file !" << std::endl;
        std::exit(1);
    }
    lap("read file");
    if (!LoadModelParallel(text.data.get(), text.size)) {
        TokenStream ts(text.data.get(), text.data.get() + text.size);
        std::vector<vec4> positions;
        std::vector<vec3> normals;
        int current_material = 0;
        const char *b, *e;
        while (ts.next(b, e)) {
            if (is_tok(b, e, "mtllib")) {
                if (!ts.next(b, e)) break;
                mtl_files_.push_back(Config::get().model_path + std::string(b, e));
                read_mtl(mtl_files_.back(), *this);
            } else if (is_tok(b, e, "v")) {
                vec4 p(1);
                p.x = ts.next_float(); p.y = ts.next_float(); p.z = ts.next_float();
                positions.push_back(p);
            } else if (is_tok(b, e, "vt")) {
                ts.next_float(); ts.next_float();
            } else if (is_tok(b, e, "vn")) {
                vec3 n;
                n.x = ts.next_float(); n.y = ts.next_float(); n.z = ts.next_float();
                normals.push_back(n);
            } else if (is_tok(b, e, "f")) {
                int idx[3][3];
                for (int i = 0; i < 3; ++i) {
                    if (!ts.next(b, e)) { b = e = ""; }
                    corner(b, e, idx[i]);
                }
                Triangle t({positions.at(idx[0][0]), positions.at(idx[1][0]), positions.at(idx[2][0])}, current_material);
                if (idx[0][2] >= 0) t.SetNormal(normals.at(idx[0][2]));
                triangles_.push_back(t);
            } else if (is_tok(b, e, "usemtl")) {
                current_material = ts.next_int();
            }
        }
    }
    if (!cache_dir.empty() && fresh) SaveCache(cache_name(cache_dir, path), path);
}

PtScene Scene::abi() const {
    const size_t n = triangles_.size(), m = materials_.size();
    vertices_.resize(9 * n); planes_.resize(4 * n); squares_.resize(n); mat_ids_.resize(n); mats_.resize(10 * m);
    mats_ext_.resize(2 * m);
    bool any_refract = false;
    auto flatten = [&](size_t i0, size_t i1) {
        for (size_t i = i0; i < i1; ++i) {
            const Triangle &t = triangles_[i];
            for (int k = 0; k < 3; ++k)
                for (int a = 0; a < 3; ++a) vertices_[9 * i + 3 * k + a] = t.GetVertices()[k][a];
            const vec4 p = t.GetPlane();
            planes_[4 * i] = p.x; planes_[4 * i + 1] = p.y; planes_[4 * i + 2] = p.z; planes_[4 * i + 3] = p.w;
            squares_[i] = t.GetSquare();
            mat_ids_[i] = t.GetMaterialId();
        }
    };
    const size_t parts = n >= 200000 ? std::min<size_t>(std::max(1u, std::thread::hardware_concurrency()), 16) : 1;
    run_threads(parts, [&](size_t k) { flatten(n / parts * k, k + 1 == parts ? n : n / parts * (k + 1)); });
    for (size_t j = 0; j < m; ++j) {
        const MaterialCharacteristics &c = materials_[j].GetCharacteristics();
        const float row[10] = {c.Ke.x, c.Ke.y, c.Ke.z, c.Kd.x, c.Kd.y, c.Kd.z, c.Ks.x, c.Ks.y, c.Ks.z, c.Ns};
        std::copy(row, row + 10, mats_.begin() + 10 * j);
        mats_ext_[2 * j] = c.Ni; mats_ext_[2 * j + 1] = c.Tr;
        any_refract = any_refract || materials_[j].HasRefract();
    }
    PtScene s{};
    s.n_triangles = static_cast<int32_t>(n);
    s.vertices = vertices_.data(); s.planes = planes_.data(); s.squares = squares_.data();
    s.material_ids = mat_ids_.data();
    s.n_materials = static_cast<int32_t>(m);
    s.materials = mats_.data();
    if (any_refract) s.materials_ext = mats_ext_.data();
    if (skybox_.width() && skybox_.height()) {
        s.sky_width = static_cast<int32_t>(skybox_.width());
        s.sky_height = static_cast<int32_t>(skybox_.height());
        s.sky_bgr = skybox_.data();
    }
    return s;
}

int Scene::Intersect(const std::vector<Ray> &rays, const PtConfig &cfg, std::vector<int32_t> &ids, std::vector<float> &t) const {
    const PtScene s = abi();
    std::vector<float> rays6(6 * rays.size());
    for (size_t i = 0; i < rays.size(); ++i) {  // pt_intersect normalises the direction exactly as Ray's constructor did (ray.h:21-25)
        const vec4 o = rays[i].GetBegin(), d = rays[i].GetGivenDirection();
        rays6[6 * i] = o.x; rays6[6 * i + 1] = o.y; rays6[6 * i + 2] = o.z;
        rays6[6 * i + 3] = d.x; rays6[6 * i + 4] = d.y; rays6[6 * i + 5] = d.z;
    }
    ids.assign(rays.size(), -1);
    t.assign(rays.size(), 0.0f);
    return pt_intersect(&s, &cfg, rays6.data(), static_cast<int64_t>(rays.size()), ids.data(), t.data());
}

int Scene::Render(const PtConfig &cfg) {
    const PtScene s = abi();
    const size_t W = cfg.width, H = cfg.height;
    sum_.assign(3 * W * H, 0.0f); sum2_.assign(3 * W * H, 0.0f);
    count_.assign(W * H, 0);
    weight_.assign(W * H, 0.0f);
    const int rc = pt_render_accum_w(&s, &cfg, sum_.data(), sum2_.data(), count_.data(), weight_.data());
    if (rc) return rc;
    for (size_t y = 0; y < H; ++y)
        for (size_t x = 0; x < W; ++x) {
            const size_t p = y * W + x;
            color_map_[x][y] = vec3(sum_[3 * p], sum_[3 * p + 1], sum_[3 * p + 2]);
            color2_map_[x][y] = vec3(sum2_[3 * p], sum2_[3 * p + 1], sum2_[3 * p + 2]);
            samples_count_[x][y] = count_[p];
        }
    rr_ = cfg.rr_depth > 0;
    return 0;
}

int Scene::Resolve(const PtPostConfig &post, std::vector<uint8_t> &rgb8, PtDispersion &stats) const {
    rgb8.assign(3 * static_cast<size_t>(post.width) * post.height, 0);
    return pt_postprocess(sum_.data(), sum2_.data(), count_.data(), rr_ ? weight_.data() : nullptr, &post, rgb8.data(), &stats);
}

int Scene::RenderImage(const PtConfig &cfg, const PtPostConfig &post, std::vector<uint8_t> &rgb8, PtDispersion &stats) {
    const PtScene s = abi();
    rgb8.assign(3 * static_cast<size_t>(post.width) * post.height, 0);
    return pt_render_image(&s, &cfg, &post, rgb8.data(), &stats);
}

int Scene::RenderProgressive(const PtConfig &cfg, const PtPostConfig &post, int update, long long time_limit_ms,
                             PtPreviewFn on_update, void *user, std::vector<uint8_t> &rgb8, PtDispersion &stats,
                             int &samples_done) {
    const PtScene s = abi();
    rgb8.assign(3 * static_cast<size_t>(post.width) * post.height, 0);
    int32_t done = 0;
    const int rc = pt_render_progressive(&s, &cfg, &post, update, time_limit_ms, on_update, user, rgb8.data(), &stats, &done);
    samples_done = done;
    return rc;
}
