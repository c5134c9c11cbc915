// obj_reader.h — TEST INFRASTRUCTURE (oracle side only).
// Minimal Wavefront OBJ reader standing in for Assimp inside the oracle builds.  It yields what mesh.cpp:53-120
// receives from Assimp's OBJ importer under aiProcess_Triangulate | aiProcess_FlipUVs: per-face-corner
// (unshared) positions / normals / uvs, polygons fan-triangulated, v flipped to 1-v, and all-zero normals / uvs
// when the file provides none (mesh.cpp:83-99).
#pragma once
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

namespace objr {

struct TriSoup {
    std::vector<float> pos, nrm, uv;   // 9 / 9 / 6 floats per triangle
    size_t n_tri() const { return pos.size() / 9; }
};

inline bool load_obj(const std::string& path, TriSoup& out)
{
    FILE* f = fopen(path.c_str(), "rb");
    if (!f) return false;
    std::vector<float> V, N, T;
    struct Corner { int v, t, n; };
    std::vector<Corner> face;
    char line[4096];
    bool any_n = false, any_t = false;
    std::vector<Corner> corners;           // 3 per triangle
    while (fgets(line, sizeof line, f)) {
        char* p = line; while (*p == ' ' || *p == '\t') ++p;
        if (p[0] == 'v' && (p[1] == ' ' || p[1] == '\t')) {
            float x = 0, y = 0, z = 0; sscanf(p + 2, "%f %f %f", &x, &y, &z); V.push_back(x); V.push_back(y); V.push_back(z);
        } else if (p[0] == 'v' && p[1] == 'n') {
            float x = 0, y = 0, z = 0; sscanf(p + 3, "%f %f %f", &x, &y, &z); N.push_back(x); N.push_back(y); N.push_back(z);
        } else if (p[0] == 'v' && p[1] == 't') {
            float x = 0, y = 0; sscanf(p + 3, "%f %f", &x, &y); T.push_back(x); T.push_back(y);
        } else if (p[0] == 'f' && (p[1] == ' ' || p[1] == '\t')) {
            face.clear();
            char* q = p + 2;
            while (*q) {
                while (*q == ' ' || *q == '\t') ++q;
                if (*q == '\n' || *q == '\r' || *q == 0) break;
                Corner c{0, 0, 0};
                c.v = (int)strtol(q, &q, 10);
                if (*q == '/') { ++q; if (*q != '/') c.t = (int)strtol(q, &q, 10); if (*q == '/') { ++q; c.n = (int)strtol(q, &q, 10); } }
                int nv = (int)(V.size() / 3), nt = (int)(T.size() / 2), nn = (int)(N.size() / 3);
                if (c.v < 0) c.v = nv + c.v + 1; if (c.t < 0) c.t = nt + c.t + 1; if (c.n < 0) c.n = nn + c.n + 1;
                face.push_back(c);
                while (*q && *q != ' ' && *q != '\t' && *q != '\n' && *q != '\r') ++q;
            }
            for (size_t k = 1; k + 1 < face.size(); ++k) { corners.push_back(face[0]); corners.push_back(face[k]); corners.push_back(face[k + 1]); }
        }
    }
    fclose(f);
    for (const Corner& c : corners) { if (c.n > 0) any_n = true; if (c.t > 0) any_t = true; }
    for (const Corner& c : corners) {
        int vi = c.v - 1;
        if (vi < 0 || (size_t)vi * 3 + 2 >= V.size() + 0) { out.pos.push_back(0); out.pos.push_back(0); out.pos.push_back(0); }
        else { out.pos.push_back(V[vi * 3]); out.pos.push_back(V[vi * 3 + 1]); out.pos.push_back(V[vi * 3 + 2]); }
        if (any_n && c.n > 0 && (size_t)(c.n - 1) * 3 + 2 < N.size()) { out.nrm.push_back(N[(c.n - 1) * 3]); out.nrm.push_back(N[(c.n - 1) * 3 + 1]); out.nrm.push_back(N[(c.n - 1) * 3 + 2]); }
        else { out.nrm.push_back(0); out.nrm.push_back(0); out.nrm.push_back(0); }
        if (any_t && c.t > 0 && (size_t)(c.t - 1) * 2 + 1 < T.size()) { out.uv.push_back(T[(c.t - 1) * 2]); out.uv.push_back(1.0f - T[(c.t - 1) * 2 + 1]); }
        else { out.uv.push_back(0); out.uv.push_back(0); }
    }
    return true;
}

} // namespace objr
