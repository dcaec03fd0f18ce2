/* shared-memory bank-conflict model of K4's gathers on the floodplain mesh: wavefronts per warp instruction for the
 * phase-2 l-side / r-side loads (index = local cell) and the phase-3 flux gathers (index = edge slot).
 * 64-bit loads: one request per half-warp; wavefronts(half) = max over the 16 bank pairs of the number of DISTINCT words. */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include "rdyb200_mesh.h"
#include "device_layout.h"

static int half_wavefronts(const int *idx, int n) { /* idx < 0: inactive lane */
  int cnt[16] = {0}, seen[16][16], w = 0;
  for (int i = 0; i < n; ++i) {
    if (idx[i] < 0) continue;
    int b = idx[i] & 15, dup = 0;
    for (int j = 0; j < cnt[b]; ++j) if (seen[b][j] == idx[i]) dup = 1;
    if (!dup) seen[b][cnt[b]++] = idx[i];
    if (cnt[b] > w) w = cnt[b];
  }
  return w;
}

int main(int argc, char **argv) {
  int nx = argc > 1 ? atoi(argv[1]) : 500;
  RdybHostMesh m;
  if (rdyb_hostmesh_generate_floodplain(nx, nx, nx * 10.0, nx * 10.0, 0.2, 1.0, 12345, 1, 1, 0, &m)) return 1;
  RdybMeshView v;
  rdyb_hostmesh_view(m, &v);
  RdybLayout L; memset(&L, 0, sizeof L);
  if (rdyb_layout_build(&v, 0, NULL, NULL, 1, &L)) { printf("layout failed\n"); return 1; }
  RdybTiledLayout T; memset(&T, 0, sizeof T);
  if (rdyb_tiled_layout_build(&L, 224, &T)) { printf("tiled failed\n"); return 1; }
  const int W = L.width; const size_t S = L.stride;
  double wl = 0, wr = 0, w3 = 0; long nl = 0, n3 = 0;
  for (int t = 0; t < T.num_tiles; ++t) {
    const int32_t *d = T.tile_desc + 8 * (size_t)t;
    int cb = d[0], nc = d[1], eo = d[2], ne = d[3];
    for (int e0 = 0; e0 < ne; e0 += 224)               /* the 224 consumer threads take edges e0 .. e0+223 */
      for (int w0 = e0; w0 < ne && w0 < e0 + 224; w0 += 32) {
        int a = 0, b = 0;
        for (int h = 0; h < 2; ++h) {
          int il[16], ir[16];
          for (int j = 0; j < 16; ++j) {
            int e = w0 + 16 * h + j;
            if (e < ne && e < e0 + 224) { uint32_t lr = T.edge_lr[eo + e]; il[j] = lr & 0xffff; ir[j] = (lr >> 16) == 0xffff ? -1 : (int)(lr >> 16); }
            else il[j] = ir[j] = -1;
          }
          a += half_wavefronts(il, 16); b += half_wavefronts(ir, 16);
        }
        wl += a; wr += b; nl++;
      }
    for (int k = 0; k < W; ++k)
      for (int w0 = 0; w0 < nc; w0 += 32) {
        int a = 0;
        for (int h = 0; h < 2; ++h) {
          int ie[16];
          for (int j = 0; j < 16; ++j) {
            int c = w0 + 16 * h + j;
            if (c < nc) { uint16_t ce = T.cell_edges[(size_t)k * S + cb + c]; int e = ce & 0x7fff; ie[j] = e > 399 ? 399 : e; } else ie[j] = -1;
          }
          a += half_wavefronts(ie, 16);
        }
        w3 += a; n3++;
      }
  }
  { double exl = 0, exr = 0, exs = 0; long nt = 0, ns = 0;
    for (int t = 0; t < T.num_interior_tiles; ++t) {
      const int32_t *d = T.tile_desc + 8 * (size_t)t; int eo = d[2], ne = d[3], G = (ne + 15) / 16, nl[16] = {0}, nr[16] = {0};
      for (int e = 0; e < ne; ++e) { uint32_t lr = T.edge_lr[eo + e]; nl[lr & 15]++; if ((lr >> 16) != 0xffff) nr[(lr >> 16) & 15]++; }
      for (int b = 0; b < 16; ++b) { if (nl[b] > G) exl += nl[b] - G; if (nr[b] > G) exr += nr[b] - G; }
      nt++;
    }
    printf("excess edges per tile over G per bank: l %.2f r %.2f\n", exl / nt, exr / nt); }
  printf("tiles %d width %d max_edges %d: wavefronts per warp instruction  l-side %.3f  r-side %.3f  phase-3 %.3f\n", T.num_tiles, W, T.max_edges, wl / nl, wr / nl, w3 / n3);
  return 0;
}
