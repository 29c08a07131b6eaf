// scene_io.cu — SURVEY.md §8f.2: scene input formats feeding build_bvh.  Host code
// only (no kernels): a Wavefront OBJ reader (v / f records, polygons fanned into
// triangles, negative = relative indices, v/vt/vn forms) and a flat binary
// triangle file.  Both produce the library's triangle layout: 12 floats per
// triangle, v0.xyz,0, v1.xyz,0, v2.xyz,0.  The reference defines no scene format
// (README.md:1-2 is all it holds), so these are the two simplest carriers.
#include <cerrno>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "../../include/gpu_raytracer.h"
#include "common.cuh"

namespace {

constexpr char kTriMagic[8] = {'R', 'T', 'T', 'R', 'I', '1', 0, 0};

// parses the vertex index of one "f" token (v, v/vt, v//vn, v/vt/vn); 0 on error
long face_index(const char* tok, char** end) {
  long v = strtol(tok, end, 10);
  while (**end == '/') {  // skip texture / normal indices
    ++*end;
    if (**end != '/' && **end != ' ' && **end != '\t' && **end != '\0' && **end != '\r' &&
        **end != '\n')
      strtol(*end, end, 10);
  }
  return v;
}

}  // namespace

extern "C" int rt_load_obj(const char* path, float** tris_out, uint64_t* n_out) {
  if (!path || !tris_out || !n_out) return rt::fail(RT_ERR_INVALID, "rt_load_obj: NULL argument");
  *tris_out = nullptr;
  *n_out = 0;
  FILE* f = fopen(path, "rb");
  if (!f) return rt::fail(RT_ERR_INVALID, "rt_load_obj: cannot open %s: %s", path, strerror(errno));
  std::vector<float> verts;   // xyz
  std::vector<float> tris;    // 12 floats each
  std::vector<long> poly;
  char* line = nullptr;
  size_t cap = 0;
  long lineno = 0;
  int rc = RT_OK;
  while (getline(&line, &cap, f) >= 0) {
    ++lineno;
    const char* p = line;
    while (*p == ' ' || *p == '\t') ++p;
    if (p[0] == 'v' && (p[1] == ' ' || p[1] == '\t')) {
      char* e = nullptr;
      float xyz[3];
      const char* q = p + 1;
      bool ok = true;
      for (int k = 0; k < 3; ++k) {
        xyz[k] = strtof(q, &e);
        if (e == q) ok = false;
        q = e;
      }
      if (!ok) {
        rc = rt::fail(RT_ERR_INVALID, "rt_load_obj: %s:%ld: malformed vertex", path, lineno);
        break;
      }
      verts.insert(verts.end(), xyz, xyz + 3);
    } else if (p[0] == 'f' && (p[1] == ' ' || p[1] == '\t')) {
      poly.clear();
      const char* q = p + 1;
      const long nv = (long)(verts.size() / 3);
      while (true) {
        while (*q == ' ' || *q == '\t') ++q;
        if (*q == '\0' || *q == '\n' || *q == '\r' || *q == '#') break;
        char* e = nullptr;
        long v = face_index(q, &e);
        if (e == q) { v = 0; }
        q = e;
        const long idx = v > 0 ? v - 1 : (v < 0 ? nv + v : -1);
        if (idx < 0 || idx >= nv) {
          rc = rt::fail(RT_ERR_INVALID, "rt_load_obj: %s:%ld: vertex index %ld out of range (%ld vertices)",
                        path, lineno, v, nv);
          break;
        }
        poly.push_back(idx);
      }
      if (rc != RT_OK) break;
      if (poly.size() < 3) {
        rc = rt::fail(RT_ERR_INVALID, "rt_load_obj: %s:%ld: face with %zu vertices", path, lineno,
                      poly.size());
        break;
      }
      for (size_t k = 1; k + 1 < poly.size(); ++k) {  // fan around the first vertex
        const long ids[3] = {poly[0], poly[k], poly[k + 1]};
        for (int c = 0; c < 3; ++c) {
          const float* v = &verts[(size_t)ids[c] * 3];
          const float rec[4] = {v[0], v[1], v[2], 0.0f};
          tris.insert(tris.end(), rec, rec + 4);
        }
      }
    }
    // every other record (vt, vn, o, g, s, usemtl, mtllib, comments) is ignored
  }
  free(line);
  fclose(f);
  if (rc != RT_OK) return rc;
  const uint64_t n = tris.size() / 12;
  if (n > RT_MAX_TRIS) return rt::fail(RT_ERR_LIMIT, "rt_load_obj: %llu triangles exceed RT_MAX_TRIS",
                                       (unsigned long long)n);
  float* out = (float*)malloc(n ? tris.size() * sizeof(float) : 4);
  if (!out) return rt::fail(RT_ERR_OOM, "rt_load_obj: host allocation failed");
  if (n) memcpy(out, tris.data(), tris.size() * sizeof(float));
  *tris_out = out;
  *n_out = n;
  return RT_OK;
}

extern "C" int rt_save_tris(const char* path, const float* tris, uint64_t n) {
  if (!path || (n && !tris)) return rt::fail(RT_ERR_INVALID, "rt_save_tris: NULL argument");
  FILE* f = fopen(path, "wb");
  if (!f) return rt::fail(RT_ERR_INVALID, "rt_save_tris: cannot open %s: %s", path, strerror(errno));
  bool ok = fwrite(kTriMagic, 1, 8, f) == 8 && fwrite(&n, 8, 1, f) == 1 &&
            (n == 0 || fwrite(tris, 48, n, f) == n);
  ok = (fclose(f) == 0) && ok;
  if (!ok) return rt::fail(RT_ERR_INVALID, "rt_save_tris: short write to %s", path);
  return RT_OK;
}

extern "C" int rt_load_tris(const char* path, float** tris_out, uint64_t* n_out) {
  if (!path || !tris_out || !n_out) return rt::fail(RT_ERR_INVALID, "rt_load_tris: NULL argument");
  *tris_out = nullptr;
  *n_out = 0;
  FILE* f = fopen(path, "rb");
  if (!f) return rt::fail(RT_ERR_INVALID, "rt_load_tris: cannot open %s: %s", path, strerror(errno));
  char magic[8];
  uint64_t n = 0;
  if (fread(magic, 1, 8, f) != 8 || memcmp(magic, kTriMagic, 8) != 0 || fread(&n, 8, 1, f) != 1) {
    fclose(f);
    return rt::fail(RT_ERR_INVALID, "rt_load_tris: %s is not a triangle file", path);
  }
  if (n > RT_MAX_TRIS) {
    fclose(f);
    return rt::fail(RT_ERR_LIMIT, "rt_load_tris: %llu triangles exceed RT_MAX_TRIS", (unsigned long long)n);
  }
  float* out = (float*)malloc(n ? n * 48 : 4);
  if (!out) {
    fclose(f);
    return rt::fail(RT_ERR_OOM, "rt_load_tris: host allocation failed");
  }
  const bool ok = n == 0 || fread(out, 48, n, f) == n;
  fclose(f);
  if (!ok) {
    free(out);
    return rt::fail(RT_ERR_INVALID, "rt_load_tris: %s is truncated", path);
  }
  *tris_out = out;
  *n_out = n;
  return RT_OK;
}

extern "C" void rt_free_host(void* p) { free(p); }
