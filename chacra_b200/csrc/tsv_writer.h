// tsv_writer.h -- the contact-list TSV of contact_consumer (gc/contact_calc/compute_contacts.py:306-359)
// written natively: the records drained from the device are ordered, de-duplicated and formatted by host threads and
// go to the file descriptor in frame order.  (The reference builds every line from Python strings in one consumer
// process; at ~5 000 lines per frame this stage, not the geometry, decides how long the drop-in script runs.)
//
//   line = frame \t itype \t atom_1 \t atom_2 [\t water_1 [\t water_2]] [\t "%.3f" distance] \n
//   atom_1 / atom_2 ordered as strings (label ranks), the waters of a two-water bridge swapped with them (:343-348);
//   identical lines of a fragment once (:350-359; the frame is part of the line, so per frame is the same thing);
//   inside a frame: interaction blocks in compute order, then ascending atom indices (the reference's order inside a
//   block is undefined).
#pragma once
#include <algorithm>
#include <chrono>
#include <cerrno>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <new>
#include <string>
#include <thread>
#include <unistd.h>
#include <unordered_set>
#include <vector>

namespace tsvw {

inline char* put_uint(char* p, unsigned long long v) {
    char tmp[24]; int n = 0;
    do { tmp[n++] = (char)('0' + v % 10); v /= 10; } while (v);
    while (n) *p++ = tmp[--n];
    return p;
}

inline char* put_int(char* p, long long v) {
    if (v < 0) { *p++ = '-'; return put_uint(p, (unsigned long long)(-(v + 1)) + 1ull); }
    return put_uint(p, (unsigned long long)v);
}

// exactly printf("%.3f", (double)f): a binary32 value is m * 2^e with a 24-bit m, so m * 1000 / 2^k is exact in 64-bit
// integers and the rounding (to nearest, ties to even, as glibc and CPython do on the exact value) can be decided exactly
inline char* put_fixed3(char* p, float f) {
    uint32_t bits; memcpy(&bits, &f, 4);
    const int exp = (int)((bits >> 23) & 0xff);
    uint64_t m = bits & 0x7fffffu;
    if (exp == 0xff || (bits >> 31)) return p + snprintf(p, 48, "%.3f", (double)f);       // inf, nan, negative: not a distance
    int e;
    if (exp == 0) e = -149; else { m |= 0x800000u; e = exp - 150; }
    uint64_t q;                                         // value * 1000, rounded
    if (e >= 0) {
        if (e > 30) return p + snprintf(p, 48, "%.3f", (double)f);
        q = (m << e) * 1000ull;
    } else {
        const int k = -e;
        if (k > 62) q = 0;                              // below 2^-38: prints as 0.000
        else {
            const uint64_t num = m * 1000ull, half = 1ull << (k - 1), rem = num & ((1ull << k) - 1ull);
            q = num >> k;
            if (rem > half || (rem == half && (q & 1ull))) ++q;
        }
    }
    p = put_uint(p, q / 1000ull);
    const unsigned r = (unsigned)(q % 1000ull);
    *p++ = '.'; *p++ = (char)('0' + r / 100); *p++ = (char)('0' + r / 10 % 10); *p++ = (char)('0' + r % 10);
    return p;
}

// growing text buffer without the zero fill of std::vector::resize
struct Text {
    char* p = nullptr; size_t n = 0, cap = 0;
    Text() = default;
    Text(const Text&) = delete;
    Text& operator=(const Text&) = delete;
    ~Text() { free(p); }
    bool room(size_t extra) {
        if (n + extra <= cap) return true;
        size_t want = std::max(cap + cap / 2, n + extra + 4096);
        char* q = static_cast<char*>(realloc(p, want));
        if (!q) return false;
        p = q; cap = want;
        return true;
    }
    const char* data() const { return p; }
    size_t size() const { return n; }
};

struct Rec { int32_t frame, itype, a1, a2, w1, w2; float dist; int32_t reserved; };

struct Tables {
    const char* label_text; const int64_t* label_off; const int32_t* label_rank; int32_t n_atoms; bool ranks_unique;
    const char* itype_text; const int32_t* itype_off; const int32_t* block_rank; int32_t n_itypes;
    long long beg, stride; bool distout;
};

inline char* put_label(char* p, const Tables& T, int a) {
    const int64_t b = T.label_off[a], n = T.label_off[a + 1] - b;
    memcpy(p, T.label_text + b, (size_t)n);
    return p + n;
}

// one frame's records [lo, hi) -> text appended to out; returns the number of lines
inline int64_t format_frame(const Tables& T, Rec* r, int64_t lo, int64_t hi, Text& out, size_t max_line) {
    std::sort(r + lo, r + hi, [&](const Rec& x, const Rec& y) {
        const int bx = T.block_rank[x.itype], by = T.block_rank[y.itype];
        if (bx != by) return bx < by;
        if (x.a1 != y.a1) return x.a1 < y.a1;
        if (x.a2 != y.a2) return x.a2 < y.a2;
        if (x.w1 != y.w1) return x.w1 < y.w1;
        if (x.w2 != y.w2) return x.w2 < y.w2;
        if (x.itype != y.itype) return x.itype < y.itype;
        uint32_t dx, dy; memcpy(&dx, &x.dist, 4); memcpy(&dy, &y.dist, 4);
        return dx < dy;
    });
    int64_t lines = 0;
    std::unordered_set<std::string> seen;             // only for topologies in which two atoms share a label
    size_t prev_at = 0, prev_len = 0; bool have_prev = false;
    for (int64_t k = lo; k < hi; ++k) {
        const Rec& c = r[k];
        const size_t at = out.n;
        if (!out.room(max_line)) return -1;
        char* p0 = out.p + at; char* p = p0;
        p = put_int(p, T.beg + (long long)c.frame * T.stride);
        *p++ = '\t';
        const int nb = T.itype_off[c.itype], nl = T.itype_off[c.itype + 1] - nb;
        memcpy(p, T.itype_text + nb, (size_t)nl); p += nl;
        *p++ = '\t'; p = put_label(p, T, c.a1);
        *p++ = '\t'; p = put_label(p, T, c.a2);
        if (c.w1 >= 0) {
            *p++ = '\t'; p = put_label(p, T, c.w1);
            if (c.w2 >= 0) { *p++ = '\t'; p = put_label(p, T, c.w2); }
        }
        if (T.distout) { *p++ = '\t'; p = put_fixed3(p, c.dist); }
        *p++ = '\n';
        const size_t len = (size_t)(p - p0);
        bool dup;
        if (T.ranks_unique) {                           // equal lines are neighbours in this order
            dup = have_prev && len == prev_len && memcmp(out.p + prev_at, p0, len) == 0;
        } else {
            dup = !seen.emplace(p0, len).second;
        }
        if (dup) continue;
        out.n = at + len;
        prev_at = at; prev_len = len; have_prev = true;
        ++lines;
    }
    return lines;
}

inline bool write_all(int fd, const char* p, size_t n) {
    while (n) {
        const ssize_t w = ::write(fd, p, n);
        if (w < 0) { if (errno == EINTR) continue; return false; }
        p += w; n -= (size_t)w;
    }
    return true;
}

// returns lines written, -1 bad argument, -2 write error, -3 out of memory
inline int64_t write_contacts(int fd, Rec* rec, int64_t n, const Tables& T, int n_threads) {
    if (n == 0) return 0;
    if (fd < 0 || !rec || n < 0 || !T.label_text || !T.label_off || !T.label_rank || !T.itype_text || !T.itype_off || !T.block_rank) return -1;
    size_t max_label = 0;
    for (int a = 0; a < T.n_atoms; ++a) max_label = std::max<size_t>(max_label, (size_t)(T.label_off[a + 1] - T.label_off[a]));
    size_t max_name = 0;
    for (int k = 0; k < T.n_itypes; ++k) max_name = std::max<size_t>(max_name, (size_t)(T.itype_off[k + 1] - T.itype_off[k]));
    const size_t max_line = 24 + max_name + 4 * max_label + 8 + 48;
    int nt = std::max(1, std::min<int>(n_threads, 64));
    auto run = [&](int count, auto&& fn) {              // fn(t) on `count` threads, the caller's included
        std::vector<std::thread> pool;
        for (int t = 1; t < count; ++t) pool.emplace_back(fn, t);
        fn(0);
        for (auto& th : pool) th.join();
    };
    const bool timing = getenv("CHACRA_TIMING") != nullptr;
    auto now = [] { return std::chrono::steady_clock::now(); };
    auto t_start = now();
    // pass 1 (blocks of records): validate, order atom_1 / atom_2 by label, frame range
    const int nb = (int)std::min<int64_t>(nt, (n + 65535) / 65536);
    std::vector<int32_t> bmin((size_t)nb), bmax((size_t)nb);
    std::vector<int> bad((size_t)nb, 0);
    auto block = [&](int t, int64_t* lo, int64_t* hi) { *lo = n * t / nb; *hi = n * (t + 1) / nb; };
    run(nb, [&](int t) {
        int64_t lo, hi; block(t, &lo, &hi);
        int32_t mn = rec[lo].frame, mx = rec[lo].frame;
        for (int64_t k = lo; k < hi; ++k) {
            Rec& c = rec[k];
            if (c.itype < 0 || c.itype >= T.n_itypes || c.a1 < 0 || c.a1 >= T.n_atoms || c.a2 < 0 || c.a2 >= T.n_atoms ||
                c.w1 >= T.n_atoms || c.w2 >= T.n_atoms || c.frame < 0) { bad[(size_t)t] = 1; return; }
            mn = std::min(mn, c.frame); mx = std::max(mx, c.frame);
            if (T.label_rank[c.a2] < T.label_rank[c.a1]) {  // "if interaction[3] < interaction[2]": swap, and the two waters with them
                std::swap(c.a1, c.a2);
                if (c.w2 >= 0) std::swap(c.w1, c.w2);
            }
        }
        bmin[(size_t)t] = mn; bmax[(size_t)t] = mx;
    });
    for (int t = 0; t < nb; ++t) if (bad[(size_t)t]) return -1;
    const int32_t fmin = *std::min_element(bmin.begin(), bmin.end()), fmax = *std::max_element(bmax.begin(), bmax.end());
    // pass 2: counting sort by frame, one histogram per block
    const int64_t nf = (int64_t)fmax - fmin + 1;
    if (nf > (1ll << 26)) return -1;                   // one drain covers a few thousand frames; refuse absurd frame spans
    std::vector<std::vector<int64_t>> hist((size_t)nb, std::vector<int64_t>((size_t)nf, 0));
    run(nb, [&](int t) {
        int64_t lo, hi; block(t, &lo, &hi);
        int64_t* h = hist[(size_t)t].data();
        for (int64_t k = lo; k < hi; ++k) ++h[rec[k].frame - fmin];
    });
    std::vector<int64_t> start((size_t)nf + 1, 0);
    {
        int64_t run_sum = 0;
        for (int64_t f = 0; f < nf; ++f) {
            start[(size_t)f] = run_sum;
            for (int t = 0; t < nb; ++t) { const int64_t c = hist[(size_t)t][(size_t)f]; hist[(size_t)t][(size_t)f] = run_sum; run_sum += c; }
        }
        start[(size_t)nf] = run_sum;
    }
    std::unique_ptr<Rec[]> sorted(new (std::nothrow) Rec[(size_t)n]);          // not value-initialised: first touched by the scatter threads
    if (!sorted) return -3;
    run(nb, [&](int t) {
        int64_t lo, hi; block(t, &lo, &hi);
        int64_t* h = hist[(size_t)t].data();
        for (int64_t k = lo; k < hi; ++k) sorted[(size_t)h[rec[k].frame - fmin]++] = rec[k];
    });
    // threads take contiguous frame ranges holding about the same number of records; the texts leave in that order
    nt = (int)std::min<int64_t>(nt, nf);
    std::vector<int64_t> cut((size_t)nt + 1, nf);
    cut[0] = 0;
    for (int t = 1; t < nt; ++t) {
        const int64_t want = n * t / nt;
        cut[(size_t)t] = std::lower_bound(start.begin(), start.end(), want) - start.begin();
        cut[(size_t)t] = std::min<int64_t>(std::max(cut[(size_t)t], cut[(size_t)t - 1]), nf);
    }
    std::vector<Text> text((size_t)nt);
    std::vector<int64_t> lines((size_t)nt, 0);
    auto work = [&](int t) {
        Text& out = text[(size_t)t];
        if (!out.room((size_t)((start[(size_t)cut[(size_t)t + 1]] - start[(size_t)cut[(size_t)t]]) * 80 + 4096))) { lines[(size_t)t] = -1; return; }
        for (int64_t f = cut[(size_t)t]; f < cut[(size_t)t + 1]; ++f) {
            const int64_t got = format_frame(T, sorted.get(), start[(size_t)f], start[(size_t)f + 1], out, max_line);
            if (got < 0) { lines[(size_t)t] = -1; return; }
            lines[(size_t)t] += got;
        }
    };
    auto t_sorted = now();
    run(nt, work);
    auto t_text = now();
    int64_t total = 0;
    for (int t = 0; t < nt; ++t) { if (lines[(size_t)t] < 0) return -3; total += lines[(size_t)t]; }
    // a regular file takes the threads' texts in parallel at their final offsets (one thread copying into the page cache
    // tops out near 1 GB/s, below what the formatting threads produce); pipes and the like get them one after the other
    const off_t base = ::lseek(fd, 0, SEEK_CUR);
    if (base >= 0 && nt > 1) {
        std::vector<off_t> at((size_t)nt + 1, base);
        for (int t = 0; t < nt; ++t) at[(size_t)t + 1] = at[(size_t)t] + (off_t)text[(size_t)t].size();
        std::vector<int> failed((size_t)nt, 0);
        run(nt, [&](int t) {
            const char* p = text[(size_t)t].data(); size_t left = text[(size_t)t].size(); off_t o = at[(size_t)t];
            while (left) {
                const ssize_t w = ::pwrite(fd, p, left, o);
                if (w < 0) { if (errno == EINTR) continue; failed[(size_t)t] = 1; return; }
                p += w; left -= (size_t)w; o += w;
            }
        });
        for (int t = 0; t < nt; ++t) if (failed[(size_t)t]) return -2;
        if (::lseek(fd, at[(size_t)nt], SEEK_SET) < 0) return -2;
    } else {
        for (int t = 0; t < nt; ++t)
            if (!write_all(fd, text[(size_t)t].data(), text[(size_t)t].size())) return -2;
    }
    if (timing) {
        auto sec = [](auto a, auto b) { return std::chrono::duration<double>(b - a).count(); };
        fprintf(stderr, "tsv: %lld records, %d threads: by frame %.3f s, sort + text %.3f s, write %.3f s\n", (long long)n, nt,
                sec(t_start, t_sorted), sec(t_sorted, t_text), sec(t_text, now()));
    }
    return total;
}

}  // namespace tsvw
