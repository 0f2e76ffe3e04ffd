// Host-run check (no GPU needed; nvcc only because the headers are CUDA headers) of the support-candidate cells of
// csrc/sirrt_tables.hpp: for every hull of a scene pack and a large set of directions -- random, axis-aligned, on cell
// boundaries and cube edges / corners, along hull vertices and face-like differences, at tiny and huge magnitudes -- the
// winner of the candidate scan (largest dot product, lowest index on ties) must be the winner of the full scan, in FP32
// and in FP64, with the arithmetic of warp_support (csrc/sirrt_device.cuh).
//   hull_cells_host <scene.pack> [n_random]      prints one JSON line; exit code 1 on any mismatch
#include <cstdio>
#include <cstdlib>
#include <fstream>
#include <random>
#include <string>
#include <vector>

#include "sirrt_device.cuh"
#include "sirrt_tables.hpp"

template <typename T> static int scan_all(const std::vector<T>& v, int n, const T* d) {
  T bv = T(-3.0e38);
  int bi = 0;
  for (int i = 0; i < n; i++) {
    const T x = v[3 * i] * d[0] + v[3 * i + 1] * d[1] + v[3 * i + 2] * d[2];
    if (x > bv) { bv = x; bi = i; }  // ascending index, strict >: lowest index wins ties
  }
  return bi;
}
template <typename T> static int scan_cell(const std::vector<T>& v, int n, const unsigned int* cells, const unsigned short* cand, const T* d) {
  const int cell = hull_cell(d);
  if (cell < 0) return scan_all(v, n, d);
  T bv = T(-3.0e38);
  int bi = 0x7fffffff;
  for (unsigned k = cells[cell]; k < cells[cell + 1]; k++) {
    const int i = cand[k];
    const T x = v[3 * i] * d[0] + v[3 * i + 1] * d[1] + v[3 * i + 2] * d[2];
    if (x > bv || (x == bv && i < bi)) { bv = x; bi = i; }
  }
  return bi == 0x7fffffff ? 0 : bi;
}

int main(int argc, char** argv) {
  if (argc < 2) { fprintf(stderr, "usage: %s scene.pack [n_random]\n", argv[0]); return 2; }
  const long n_random = argc > 2 ? atol(argv[2]) : 200000;
  std::ifstream f(argv[1], std::ios::binary);
  std::vector<unsigned char> pack((std::istreambuf_iterator<char>(f)), std::istreambuf_iterator<char>());
  sirrt::HScene sc;
  std::string err;
  if (!sirrt::parse_pack(pack.data(), pack.size(), sc, err)) { fprintf(stderr, "pack: %s\n", err.c_str()); return 2; }
  long hulls = 0, dirs = 0, bad32 = 0, bad64 = 0, cand_total = 0, cand_max = 0;
  std::mt19937_64 rng(99);
  std::normal_distribution<double> gauss(0.0, 1.0);
  std::vector<const sirrt::HShape*> shapes;
  for (const auto& s : sc.robot.shapes) shapes.push_back(&s);
  for (const auto& ro : sc.robs) for (const auto& s : ro.robot.shapes) shapes.push_back(&s);
  for (const sirrt::HShape* hs : shapes) {
    if (hs->type != SH_HULL || hs->verts.size() / 3 <= 64) continue;
    hulls++;
    const int n = (int)(hs->verts.size() / 3);
    std::vector<float> vf(hs->verts.begin(), hs->verts.end());
    std::vector<double> vd(hs->verts.begin(), hs->verts.end());
    std::vector<unsigned int> cells;
    std::vector<unsigned short> cand;
    sirrt::build_hull_cells(hs->verts, cells, cand);
    cand_total += cells.back();
    for (int c = 0; c < SIRRT_HULL_CELLS; c++) cand_max = std::max<long>(cand_max, cells[c + 1] - cells[c]);
    auto check = [&](const double* d) {
      const float df[3] = {(float)d[0], (float)d[1], (float)d[2]};
      dirs++;
      if (scan_cell(vf, n, cells.data(), cand.data(), df) != scan_all(vf, n, df)) bad32++;
      if (scan_cell(vd, n, cells.data(), cand.data(), d) != scan_all(vd, n, d)) bad64++;
    };
    for (long i = 0; i < n_random; i++) {  // random directions at magnitudes GJK meets (1e-6 .. 1e+1)
      const double m = std::pow(10.0, -6.0 + 7.0 * (double)(rng() % 1000) / 999.0);
      const double d[3] = {gauss(rng) * m, gauss(rng) * m, gauss(rng) * m};
      check(d);
    }
    const int K = SIRRT_HULL_K;
    for (int a = 0; a < 3; a++)
      for (int sg = -1; sg <= 1; sg += 2)
        for (int iu = 0; iu <= K; iu++)
          for (int iv = 0; iv <= K; iv++)
            for (int jit = -1; jit <= 1; jit++) {  // cell corners / edges (and a hair to either side), cube edges and corners
              double d[3];
              d[a] = sg;
              d[(a + 1) % 3] = -1.0 + 2.0 * iu / K + jit * 1e-7;
              d[(a + 2) % 3] = -1.0 + 2.0 * iv / K - jit * 3e-8;
              check(d);
            }
    for (int i = 0; i < n; i++) {  // towards the vertices themselves and along edges between index neighbours (coplanar ties)
      const double d[3] = {vd[3 * i], vd[3 * i + 1], vd[3 * i + 2]};
      check(d);
      const int j = (i + 1) % n;
      const double e[3] = {vd[3 * j] - vd[3 * i], vd[3 * j + 1] - vd[3 * i + 1], vd[3 * j + 2] - vd[3 * i + 2]};
      check(e);
      const double c[3] = {d[1] * e[2] - d[2] * e[1], d[2] * e[0] - d[0] * e[2], d[0] * e[1] - d[1] * e[0]};
      check(c);
    }
    const double ax[6][3] = {{1, 0, 0}, {-1, 0, 0}, {0, 1, 0}, {0, -1, 0}, {0, 0, 1}, {0, 0, -1}};
    for (auto& d : ax) check(d);
  }
  printf("{\"hulls\": %ld, \"directions\": %ld, \"mismatch_fp32\": %ld, \"mismatch_fp64\": %ld, \"mean_candidates_per_cell\": %.2f, \"max_candidates\": %ld}\n",
         hulls, dirs, bad32, bad64, hulls ? (double)cand_total / (hulls * SIRRT_HULL_CELLS) : 0.0, cand_max);
  return (bad32 || bad64 || !hulls) ? 1 : 0;
}
