// obj_loader.h — Wavefront OBJ reader with the reference's entry points.
//
// parseObjByName keeps the contract of CudaTest/src/Loader/obj_loader.h:20-111:
// the caller pre-allocates `points` and `idxVertex`; only "v x y z" and
// "f a b c" lines are read; face indices are stored as 1-based FLOATS, not
// decremented (obj_loader.h:62-64); afterwards the points are
// mean-centred (obj_loader.h:91-92,120-133). add_mesh(..., index_base = 1) in
// scene.h subtracts the 1.
//
// Two things the reference leaves open are closed here without changing what
// it returns for the bundled objects/small_bunny.obj:
//   - the caller can pass the capacity of its arrays; lines beyond it are
//     counted but not stored (the reference writes past the end);
//   - faces written "a//n" or "a/t/n" (objects/cbox_smallbox.obj) read the
//     vertex index before the first '/', which is what the reference's unused
//     parseFaceIndex (obj_loader.h:113-118) does, kept 1-based like the rest.
// loadObj() is the allocating convenience wrapper the CLI uses.
#pragma once
#include <cstdlib>
#include <fstream>
#include <iostream>
#include <sstream>
#include <string>
#include <utility>
#include <vector>

#include "scene.h"

inline vec3 computeMean(vec3* points, int np)                         // obj_loader.h:120-126
{
    vec3 mean(0, 0, 0);
    for (int i = 0; i < np; i++) mean += points[i];
    return mean / float(np);
}

inline void centering(vec3* points, vec3& mean, int np)               // obj_loader.h:129-133
{
    for (int i = 0; i < np; i++) points[i] -= mean;
}

inline std::pair<int, int> parseFaceIndex(std::string str)            // obj_loader.h:113-118 (0-based pair)
{
    const std::size_t pos = str.find("//");
    const std::string w1 = str.substr(0, pos);
    const std::string w2 = pos == std::string::npos ? std::string("0") : str.substr(pos + 2);
    return std::make_pair(std::atoi(w1.c_str()) - 1, std::atoi(w2.c_str()) - 1);
}

inline void parseObjByName(std::string fn, vec3* points, vec3* idxVertex, int& nPoints, int& nTriangles,
                           int maxPoints = 0x7fffffff, int maxTriangles = 0x7fffffff)
{
    std::ifstream in(fn.c_str());
    int np = 0, nt = 0;
    if (!in.is_open()) {
        std::cerr << "Can't open the file " << fn << std::endl;       // obj_loader.h:85-87: report and carry on
    } else {
        std::string line, label;
        while (std::getline(in, line)) {
            std::stringstream ss(line);
            label.clear();
            ss >> label;
            if (label == "v") {
                vec3 p(0.f);
                ss >> p[0] >> p[1] >> p[2];
                if (np < maxPoints) points[np] = p;
                np++;
            } else if (label == "f") {
                vec3 idx(0.f);
                for (int k = 0; k < 3; k++) {
                    std::string tok;
                    ss >> tok;
                    idx[k] = (float)std::atof(tok.substr(0, tok.find('/')).c_str());
                }
                if (nt < maxTriangles) idxVertex[nt] = idx;
                nt++;
            }
        }
    }
    nPoints = np < maxPoints ? np : maxPoints;
    nTriangles = nt < maxTriangles ? nt : maxTriangles;
    if (nPoints > 0) {
        vec3 mean = computeMean(points, nPoints);
        centering(points, mean, nPoints);
    }
}

// Counts the v / f lines, allocates, parses. The arrays live in `store`.
struct ObjMesh {
    std::vector<vec3> points, idx;
    MeshData mesh;                                                    // views into the vectors above
};

inline bool loadObj(const std::string& fn, ObjMesh& out)
{
    std::ifstream in(fn.c_str());
    if (!in.is_open()) return false;
    int nv = 0, nf = 0;
    std::string line;
    while (std::getline(in, line)) {
        if (line.size() > 1 && line[0] == 'v' && (line[1] == ' ' || line[1] == '\t')) nv++;
        else if (line.size() > 1 && line[0] == 'f' && (line[1] == ' ' || line[1] == '\t')) nf++;
    }
    in.close();
    out.points.assign((size_t)(nv > 0 ? nv : 1), vec3(0.f));
    out.idx.assign((size_t)(nf > 0 ? nf : 1), vec3(0.f));
    parseObjByName(fn, out.points.data(), out.idx.data(), out.mesh.nPoints, out.mesh.nTriangles, nv, nf);
    out.mesh.points = out.points.data();
    out.mesh.idxVertex = out.idx.data();
    out.mesh.normals = nullptr;
    return true;
}
