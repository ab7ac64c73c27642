// tsv_counts.h -- contact-list TSV -> residue-pair frame counts on host threads: the reduction of
// get_contact_frequencies.py (gc/get_contact_frequencies.py:58-134) through res_contacts_xl / res_contacts / gen_counts
// (gc/contact_calc/transformations.py:70-150, 262-304, 446-470).
//
// Per data line: frame = int(first TAB field), 4..7 fields, itype filter on field 2, residue = first three ':' fields of
// the atom labels in fields 3 and 4, pair ordered as strings; count = number of distinct frames per pair.
// `shipped` reproduces the reader the reference actually ships (transformations.py:114-148): it starts with
// current_frame = 0, throws away the line on which the frame number changes and never flushes the last frame.
// The file is memory-mapped and cut at frame boundaries into one piece per thread (a frame's lines are consecutive in
// files written by get_dynamic_contacts); a file whose frame numbers are not strictly increasing is reported back
// (the caller then takes the line-by-line path).
#pragma once
#include <algorithm>
#include <cstdint>
#include <cstring>
#include <fcntl.h>
#include <string>
#include <sys/mman.h>
#include <sys/stat.h>
#include <thread>
#include <unistd.h>
#include <unordered_map>
#include <vector>

namespace tsvc {

struct Result {
    int64_t total_frames = 0;
    std::vector<std::string> residues;
    std::vector<int32_t> r1, r2;
    std::vector<int64_t> count;
    std::string error;
};

enum { OK = 0, E_OPEN = -1, E_PARSE = -2, E_ORDER = -3 };

struct View { const char* p; size_t n; };

inline bool is_space(char c) { return c == ' ' || c == '\t' || c == '\n' || c == '\r' || c == '\v' || c == '\f'; }

inline View strip(const char* b, const char* e) {
    while (b < e && is_space(*b)) ++b;
    while (e > b && is_space(e[-1])) --e;
    return View{b, (size_t)(e - b)};
}

// int(token) the way Python parses it here: optional sign, decimal digits (surrounding blanks were stripped with the line)
inline bool parse_int(const char* p, size_t n, long long* out) {
    size_t k = 0; bool neg = false;
    while (k < n && is_space(p[k])) ++k;
    size_t e = n;
    while (e > k && is_space(p[e - 1])) --e;
    if (k < e && (p[k] == '+' || p[k] == '-')) { neg = p[k] == '-'; ++k; }
    if (k >= e) return false;
    long long v = 0;
    for (; k < e; ++k) {
        if (p[k] == '_') { if (k + 1 >= e || p[k + 1] == '_' ) return false; continue; }
        if (p[k] < '0' || p[k] > '9') return false;
        v = v * 10 + (p[k] - '0');
    }
    *out = neg ? -v : v;
    return true;
}

inline View residue_of(View label) {                    // ":".join(label.split(":")[0:3])
    int colons = 0;
    for (size_t k = 0; k < label.n; ++k)
        if (label.p[k] == ':' && ++colons == 3) return View{label.p, k};
    return label;
}

struct Piece {
    std::unordered_map<std::string, int32_t> ids;
    std::vector<std::string> names;
    std::unordered_map<uint64_t, std::pair<int64_t, long long>> pairs;   // (id1, id2) -> (count, last frame counted)
    long long first_frame = 0, last_frame = 0; bool any = false;
    int status = OK; std::string error;
    int32_t intern(View v) {
        std::string s(v.p, v.n);
        auto it = ids.find(s);
        if (it != ids.end()) return it->second;
        const int32_t id = (int32_t)names.size();
        names.push_back(s);
        ids.emplace(std::move(s), id);
        return id;
    }
};

// the data line (not blank, not a comment) at or after p; sets frame; returns nullptr at end
inline const char* next_data_line(const char* p, const char* end, View* line, const char** after) {
    while (p < end) {
        const char* nl = static_cast<const char*>(memchr(p, '\n', (size_t)(end - p)));
        const char* le = nl ? nl : end;
        View v = strip(p, le);
        const char* nxt = nl ? nl + 1 : end;
        if (v.n && v.p[0] != '#') { *line = v; *after = nxt; return p; }
        p = nxt;
    }
    return nullptr;
}

inline bool frame_of(View line, long long* f) {
    const char* tab = static_cast<const char*>(memchr(line.p, '\t', line.n));
    return parse_int(line.p, tab ? (size_t)(tab - line.p) : line.n, f);
}

// lines [b, e) are whole frame groups; prev_frame = frame of the data line before b (0 at the start of the file)
inline void reduce_piece(const char* b, const char* e, long long prev_frame, bool shipped, bool is_last_piece,
                         const std::vector<std::string>& keep, Piece& P) {
    const char* p = b;
    View line; const char* after;
    long long cur = prev_frame;
    std::vector<std::pair<int32_t, int32_t>> pending;     // the open frame's pairs (shipped: flushed when the frame changes)
    auto flush = [&](long long frame) {
        for (auto& pr : pending) {
            const uint64_t key = ((uint64_t)(uint32_t)pr.first << 32) | (uint32_t)pr.second;
            auto it = P.pairs.find(key);
            if (it == P.pairs.end()) P.pairs.emplace(key, std::make_pair((int64_t)1, frame));
            else if (it->second.second != frame) { ++it->second.first; it->second.second = frame; }
        }
        pending.clear();
    };
    bool open = false;                                    // a frame group of this piece is open
    while (next_data_line(p, e, &line, &after)) {
        p = after;
        // fields
        View tok[8]; int nt = 0;
        const char* q = line.p; const char* le = line.p + line.n;
        while (nt < 8) {
            const char* tab = static_cast<const char*>(memchr(q, '\t', (size_t)(le - q)));
            tok[nt++] = View{q, (size_t)((tab ? tab : le) - q)};
            if (!tab) break;
            q = tab + 1;
        }
        long long frame;
        if (!parse_int(tok[0].p, tok[0].n, &frame)) { P.status = E_PARSE; P.error = "First column isn't a integer"; return; }
        if (nt < 4 || nt > 7) { P.status = E_PARSE; P.error = "Invalid number of tokens"; return; }
        const bool changed = frame != cur;
        if (changed) {
            if (open) flush(cur);
            if (P.any && frame <= P.last_frame) { P.status = E_ORDER; return; }
            if (!P.any) { P.first_frame = frame; P.any = true; }
            P.last_frame = frame;
            cur = frame; open = true;
            if (shipped) continue;                        // the line on which the frame number changes is lost
        } else if (!open) {                               // lines of frame `prev_frame` at the very start (frame 0 of the file)
            if (P.any && frame <= P.last_frame) { P.status = E_ORDER; return; }
            if (!P.any) { P.first_frame = frame; P.any = true; }
            P.last_frame = frame;
            open = true;
        }
        if (!keep.empty()) {
            bool ok = false;
            for (const std::string& k : keep) if (k.size() == tok[1].n && memcmp(k.data(), tok[1].p, tok[1].n) == 0) { ok = true; break; }
            if (!ok) continue;
        }
        View a = residue_of(tok[2]), c = residue_of(tok[3]);
        const size_t m = std::min(a.n, c.n);
        const int cmp = memcmp(c.p, a.p, m);
        if (cmp < 0 || (cmp == 0 && c.n < a.n)) std::swap(a, c);     // (r2, r1) if r2 < r1
        pending.emplace_back(P.intern(a), P.intern(c));
    }
    if (open && !(shipped && is_last_piece)) flush(cur);  // the reference never flushes what is pending at end of file
}

inline int count_file(const char* path, const std::vector<std::string>& keep, bool shipped, int n_threads, Result& R) {
    const int fd = ::open(path, O_RDONLY);
    if (fd < 0) { R.error = std::string("cannot open ") + path; return E_OPEN; }
    struct stat st;
    if (fstat(fd, &st) != 0) { ::close(fd); R.error = "fstat failed"; return E_OPEN; }
    const size_t size = (size_t)st.st_size;
    if (size == 0) { ::close(fd); return OK; }
    const char* base = static_cast<const char*>(mmap(nullptr, size, PROT_READ, MAP_PRIVATE, fd, 0));
    ::close(fd);
    if (base == MAP_FAILED) { R.error = "mmap failed"; return E_OPEN; }
    const char* end = base + size;
    // header: every line may carry "total_frames" (transformations.py:121-123); in practice the comment lines at the top
    {
        const char* p = base;
        while (p < end) {
            const char* nl = static_cast<const char*>(memchr(p, '\n', (size_t)(end - p)));
            const char* le = nl ? nl : end;
            View v = strip(p, le);
            if (v.n && v.p[0] != '#') break;
            const std::string s(v.p, v.n);
            if (s.find("total_frames") != std::string::npos) {
                const size_t sp = s.find(' ');
                if (sp != std::string::npos) {
                    size_t e2 = s.find(' ', sp + 1);
                    const std::string second = s.substr(sp + 1, e2 == std::string::npos ? std::string::npos : e2 - sp - 1);
                    long long v2;
                    const size_t colon = second.find(':');
                    const std::string num = second.substr(colon == std::string::npos ? 0 : colon + 1);
                    if (parse_int(num.data(), num.size(), &v2)) R.total_frames = v2;
                    else { munmap(const_cast<char*>(base), size); R.error = "bad total_frames header"; return E_PARSE; }
                }
            }
            if (!nl) break;
            p = nl + 1;
        }
    }
    // cut into pieces at frame boundaries
    int nt = std::max(1, std::min(n_threads, 64));
    nt = (int)std::min<size_t>((size_t)nt, size / (1u << 20) + 1);
    std::vector<const char*> cut((size_t)nt + 1, end);
    std::vector<long long> prev((size_t)nt, 0);
    cut[0] = base;
    for (int t = 1; t < nt; ++t) {
        const char* p = base + size / (size_t)nt * (size_t)t;
        p = static_cast<const char*>(memchr(p, '\n', (size_t)(end - p)));
        if (!p) { cut[(size_t)t] = end; continue; }
        ++p;
        // frame of the last data line before p, then advance to the first line with another frame
        const char* q = p; long long before = 0; bool have = false;
        while (q > base && !have) {
            const char* le = q - 1;                       // the '\n' ending the previous line
            const char* lb = le;
            while (lb > base && lb[-1] != '\n') --lb;
            View v = strip(lb, le);
            if (v.n && v.p[0] != '#') { have = frame_of(v, &before); if (!have) break; }
            q = lb;
        }
        if (!have) { cut[(size_t)t] = cut[(size_t)t - 1]; prev[(size_t)t] = prev[(size_t)t - 1]; continue; }
        View line; const char* after; long long f = before;
        while (next_data_line(p, end, &line, &after)) {
            if (!frame_of(line, &f) || f != before) break;
            p = after;
        }
        const char* at = next_data_line(p, end, &line, &after) ? line.p : end;
        // back up to the start of that line (strip may have skipped leading blanks)
        while (at > base && at < end && at[-1] != '\n') --at;
        cut[(size_t)t] = std::max(at, cut[(size_t)t - 1]);
        prev[(size_t)t] = before;
    }
    std::vector<Piece> pieces((size_t)nt);
    int last_nonempty = 0;
    for (int t = 0; t < nt; ++t) if (cut[(size_t)t] < cut[(size_t)t + 1]) last_nonempty = t;
    {
        std::vector<std::thread> pool;
        auto work = [&](int t) {
            if (cut[(size_t)t] < cut[(size_t)t + 1])
                reduce_piece(cut[(size_t)t], cut[(size_t)t + 1], prev[(size_t)t], shipped, t == last_nonempty, keep, pieces[(size_t)t]);
        };
        for (int t = 1; t < nt; ++t) pool.emplace_back(work, t);
        work(0);
        for (auto& th : pool) th.join();
    }
    munmap(const_cast<char*>(base), size);
    long long last = 0; bool any = false;
    for (int t = 0; t < nt; ++t) {
        Piece& P = pieces[(size_t)t];
        if (P.status != OK) { R.error = P.error; return P.status; }
        if (!P.any) continue;
        if (any && P.first_frame <= last) return E_ORDER;
        last = P.last_frame; any = true;
    }
    // merge: residues by name, pairs by merged ids
    std::unordered_map<std::string, int32_t> ids;
    std::unordered_map<uint64_t, int64_t> acc;
    for (int t = 0; t < nt; ++t) {
        Piece& P = pieces[(size_t)t];
        std::vector<int32_t> map(P.names.size());
        for (size_t k = 0; k < P.names.size(); ++k) {
            auto it = ids.find(P.names[k]);
            if (it == ids.end()) { it = ids.emplace(P.names[k], (int32_t)R.residues.size()).first; R.residues.push_back(P.names[k]); }
            map[k] = it->second;
        }
        for (auto& kv : P.pairs) {
            const uint64_t key = ((uint64_t)(uint32_t)map[(size_t)(kv.first >> 32)] << 32) | (uint32_t)map[(size_t)(kv.first & 0xffffffffu)];
            acc[key] += kv.second.first;
        }
    }
    R.r1.reserve(acc.size()); R.r2.reserve(acc.size()); R.count.reserve(acc.size());
    for (auto& kv : acc) { R.r1.push_back((int32_t)(kv.first >> 32)); R.r2.push_back((int32_t)(kv.first & 0xffffffffu)); R.count.push_back(kv.second); }
    return OK;
}

}  // namespace tsvc
