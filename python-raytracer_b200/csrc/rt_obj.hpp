// rt_obj.hpp - Wavefront OBJ reader for large meshes (host code): the `v` / `vn` / `f` rows igl.read_obj hands to
// Mesh.__init__ (mesh.py:20-26).  Same conventions as python-raytracer_b200/geometry.py:read_obj: coordinates parsed as
// doubles (correctly rounded, like Python's float()), the first index of each `a/b/c` corner token, 1-based or
// negative (relative to the vertices read so far), polygons fan-triangulated.  A row that does not parse is an error
// (the Python reader raises there), reported with its line number.
#pragma once
#include <cerrno>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

struct ObjData {
    std::vector<double> verts, norms;
    std::vector<int32_t> faces;
};

namespace objio {

inline bool is_space(char c) { return c == ' ' || c == '\t' || c == '\r' || c == '\n' || c == '\v' || c == '\f'; }

inline bool parse3(const char* p, const char* end, std::vector<double>& out) {
    for (int k = 0; k < 3; k++) {
        while (p < end && is_space(*p)) p++;
        if (p >= end) return false;
        char* q = nullptr;
        const double v = strtod(p, &q);
        if (q == p || (q < end && !is_space(*q))) return false;
        out.push_back(v);
        p = q;
    }
    return true;
}

// returns "" on success, else the error text
inline std::string load(const char* path, ObjData& d) {
    FILE* fh = fopen(path, "rb");
    if (!fh) return std::string("cannot open ") + path + ": " + strerror(errno);
    std::string buf;
    {
        fseek(fh, 0, SEEK_END);
        const long sz = ftell(fh);
        fseek(fh, 0, SEEK_SET);
        buf.resize(sz > 0 ? (size_t)sz : 0);
        const size_t got = buf.empty() ? 0 : fread(&buf[0], 1, buf.size(), fh);
        fclose(fh);
        if (got != buf.size()) return std::string("short read on ") + path;
    }
    buf.push_back('\n');                    // sentinel: every row ends in a newline and strtod/strtol stop there
    const char* p = buf.data();
    const char* const end_all = p + buf.size();
    long line_no = 0;
    std::vector<int32_t> idx;
    while (p < end_all) {
        const char* eol = (const char*)memchr(p, '\n', (size_t)(end_all - p));
        if (!eol) eol = end_all;
        line_no++;
        const char* s = p;
        while (s < eol && is_space(*s)) s++;
        const char* tok_end = s;
        while (tok_end < eol && !is_space(*tok_end)) tok_end++;
        const size_t tl = (size_t)(tok_end - s);
        auto bad = [&]() { return std::string(path) + ":" + std::to_string(line_no) + ": malformed row"; };
        if (tl == 1 && s[0] == 'v') {
            if (!parse3(tok_end, eol, d.verts)) return bad();
        } else if (tl == 2 && s[0] == 'v' && s[1] == 'n') {
            if (!parse3(tok_end, eol, d.norms)) return bad();
        } else if (tl == 1 && s[0] == 'f') {
            idx.clear();
            const char* q = tok_end;
            const long nv = (long)(d.verts.size() / 3);
            for (;;) {
                while (q < eol && is_space(*q)) q++;
                if (q >= eol) break;
                char* r = nullptr;
                const long i = strtol(q, &r, 10);
                if (r == q || (r < eol && !is_space(*r) && *r != '/')) return bad();
                idx.push_back((int32_t)(i > 0 ? i - 1 : nv + i));
                q = r;
                while (q < eol && !is_space(*q)) q++;     // skip the /vt/vn part of the corner
            }
            for (size_t k = 1; k + 1 < idx.size(); k++) { d.faces.push_back(idx[0]); d.faces.push_back(idx[k]); d.faces.push_back(idx[k + 1]); }
        }
        p = eol + 1;
    }
    return "";
}

}  // namespace objio
