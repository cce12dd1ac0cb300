// Isosurface extraction from the decoded SDF grid (the step after
// reconstruct/optimizer.py:296).  The reference calls scikit-image's Lewiner marching cubes on the CPU
// (reconstruct/utils.py:120-141).  Default here: MARCHING CUBES with one vertex on every grid edge that has a
// sign change (linear interpolation) - the vertex set of the reference's mesh, up to the rare interior vertices
// Lewiner's tables add - and a case table derived by gen_mc_table.py (scikit-image is absent, its tables are not
// restated: the triangulation is unpinned).  QSP_FLAG_MESH_TETS selects the round-1 extractor instead: marching
// TETRAHEDRA on the Kuhn subdivision of every cell (6 tetrahedra around the 000-111 diagonal), which also puts
// vertices on face and body diagonals (2-3x the triangles).  Both: unique vertices, watertight, faces oriented with
// the normal towards increasing SDF.
// Output contract as the reference: vertices (V,3) float32 = index * spacing + origin, faces (F,3) int32.
#pragma once
#include "common.cuh"
#include "mc_table.inc"

#define MESH_EDGES_CUBES 0x0Bu     // edge types 0, 1, 3: the three axis edges of a grid vertex
#define MESH_EDGES_TETS 0x7Fu      // + face diagonals and the body diagonal

// edge type t (0..6) joins a grid vertex with the vertex at offset bits (t+1) = (dx | dy<<1 | dz<<2),
// where x is the slowest array axis (index = (x*D + y)*D + z)
__device__ __forceinline__ int mt_lin(int x, int y, int z, int D) { return (x * D + y) * D + z; }

__global__ void k_mt_vertex_flags(const float* __restrict__ sdf, int D, float level, unsigned edge_types,
                                  unsigned char* __restrict__ vmask, int* __restrict__ vcount) {
  int v = blockIdx.x * blockDim.x + threadIdx.x;
  if (v >= D * D * D) return;
  int z = v % D, y = (v / D) % D, x = v / (D * D);
  bool in0 = sdf[v] < level;
  unsigned m = 0;
  for (int t = 0; t < 7; ++t) {
    int b = t + 1, nx = x + (b & 1), ny = y + ((b >> 1) & 1), nz = z + ((b >> 2) & 1);
    if (((edge_types >> t) & 1u) && nx < D && ny < D && nz < D) {
      bool in1 = sdf[mt_lin(nx, ny, nz, D)] < level;
      if (in0 != in1) m |= 1u << t;
    }
  }
  vmask[v] = (unsigned char)m;
  vcount[v] = __popc(m);
}

__device__ __forceinline__ void mt_edge_point(const float* __restrict__ sdf, int D, float level, int x, int y,
                                              int z, int t, float spacing, float ox, float oy, float oz,
                                              float& px, float& py, float& pz) {
  int b = t + 1, dx = b & 1, dy = (b >> 1) & 1, dz = (b >> 2) & 1;
  float sa = sdf[mt_lin(x, y, z, D)] - level, sb = sdf[mt_lin(x + dx, y + dy, z + dz, D)] - level;
  float w = sa / (sa - sb);
  px = ((float)x + w * dx) * spacing + ox;
  py = ((float)y + w * dy) * spacing + oy;
  pz = ((float)z + w * dz) * spacing + oz;
}

__global__ void k_mt_vertices(const float* __restrict__ sdf, int D, float level, float spacing, float ox,
                              float oy, float oz, const unsigned char* __restrict__ vmask,
                              const int* __restrict__ vbase, float* __restrict__ verts) {
  int v = blockIdx.x * blockDim.x + threadIdx.x;
  if (v >= D * D * D) return;
  unsigned m = vmask[v];
  if (!m) return;
  int z = v % D, y = (v / D) % D, x = v / (D * D);
  int idx = vbase[v];
  for (int t = 0; t < 7; ++t)
    if (m & (1u << t)) {
      float px, py, pz;
      mt_edge_point(sdf, D, level, x, y, z, t, spacing, ox, oy, oz, px, py, pz);
      verts[idx * 3] = px; verts[idx * 3 + 1] = py; verts[idx * 3 + 2] = pz;
      ++idx;
    }
}

// the 6 Kuhn tetrahedra: corner bit paths 0 -> a -> a|b -> 7 for the axis permutations (a,b,c)
__constant__ unsigned char c_mt_tets[6][4] = {{0, 1, 3, 7}, {0, 1, 5, 7}, {0, 2, 3, 7},
                                             {0, 2, 6, 7}, {0, 4, 5, 7}, {0, 4, 6, 7}};

template <bool WRITE>
__global__ void k_mt_cells(const float* __restrict__ sdf, int D, float level,
                           const unsigned char* __restrict__ vmask, const int* __restrict__ vbase,
                           int* __restrict__ tcount, const int* __restrict__ tbase, int* __restrict__ faces) {
  int c = blockIdx.x * blockDim.x + threadIdx.x;
  const int C = D - 1;
  if (c >= C * C * C) return;
  int z = c % C, y = (c / C) % C, x = c / (C * C);
  float s[8];
  unsigned inside = 0;
  for (int b = 0; b < 8; ++b) {
    s[b] = sdf[mt_lin(x + (b & 1), y + ((b >> 1) & 1), z + ((b >> 2) & 1), D)] - level;
    if (s[b] < 0.f) inside |= 1u << b;
  }
  int n = 0;
  int out = WRITE ? tbase[c] : 0;
  if (inside != 0 && inside != 255) {
    // vertex index of the crossing on the edge between cube corners a < b (component-wise)
    auto eidx = [&](int a, int b) -> int {
      if (a > b) { int t = a; a = b; b = t; }
      int v = mt_lin(x + (a & 1), y + ((a >> 1) & 1), z + ((a >> 2) & 1), D);
      int t = (b ^ a) - 1;                              // b = a | diff on a Kuhn path
      return vbase[v] + __popc(vmask[v] & ((1u << t) - 1u));
    };
    auto pos = [&](int a, int b, float* p) {
      if (a > b) { int t = a; a = b; b = t; }
      float w = s[a] / (s[a] - s[b]);
      int d = b ^ a;
      p[0] = (a & 1) + w * (d & 1); p[1] = ((a >> 1) & 1) + w * ((d >> 1) & 1); p[2] = ((a >> 2) & 1) + w * ((d >> 2) & 1);
    };
    for (int k = 0; k < 6; ++k) {
      int q[4], in_cnt = 0;
      unsigned tm = 0;
      for (int u = 0; u < 4; ++u) { q[u] = c_mt_tets[k][u]; if (inside & (1u << q[u])) { tm |= 1u << u; ++in_cnt; } }
      if (in_cnt == 0 || in_cnt == 4) continue;
      int tri[2][3][2];                                 // [triangle][corner] = (a, b) cube corners of the edge
      int ntri;
      if (in_cnt == 1 || in_cnt == 3) {
        unsigned single = in_cnt == 1 ? tm : (~tm & 15u);
        int a = __ffs(single) - 1, o = 0;
        for (int u = 0; u < 4; ++u) if (u != a) { tri[0][o][0] = q[a]; tri[0][o][1] = q[u]; ++o; }
        ntri = 1;
      } else {
        int ia[2], oa[2], ni = 0, no = 0;
        for (int u = 0; u < 4; ++u) { if (tm & (1u << u)) ia[ni++] = q[u]; else oa[no++] = q[u]; }
        int e[4][2] = {{ia[0], oa[0]}, {ia[0], oa[1]}, {ia[1], oa[1]}, {ia[1], oa[0]}};
        for (int u = 0; u < 2; ++u) {
          tri[0][0][u] = e[0][u]; tri[0][1][u] = e[1][u]; tri[0][2][u] = e[2][u];
          tri[1][0][u] = e[0][u]; tri[1][1][u] = e[2][u]; tri[1][2][u] = e[3][u];
        }
        ntri = 2;
      }
      if (WRITE) {
        // centroid of the inside corners of the tetrahedron, in cell coordinates
        float ci[3] = {0.f, 0.f, 0.f};
        for (int u = 0; u < 4; ++u) if (tm & (1u << u)) { ci[0] += q[u] & 1; ci[1] += (q[u] >> 1) & 1; ci[2] += (q[u] >> 2) & 1; }
        ci[0] /= in_cnt; ci[1] /= in_cnt; ci[2] /= in_cnt;
        for (int t = 0; t < ntri; ++t) {
          float p0[3], p1[3], p2[3];
          pos(tri[t][0][0], tri[t][0][1], p0); pos(tri[t][1][0], tri[t][1][1], p1); pos(tri[t][2][0], tri[t][2][1], p2);
          float ux = p1[0] - p0[0], uy = p1[1] - p0[1], uz = p1[2] - p0[2];
          float vx = p2[0] - p0[0], vy = p2[1] - p0[1], vz = p2[2] - p0[2];
          float nx = uy * vz - uz * vy, ny = uz * vx - ux * vz, nz = ux * vy - uy * vx;
          float d = nx * (p0[0] - ci[0]) + ny * (p0[1] - ci[1]) + nz * (p0[2] - ci[2]);
          int i0 = eidx(tri[t][0][0], tri[t][0][1]), i1 = eidx(tri[t][1][0], tri[t][1][1]), i2 = eidx(tri[t][2][0], tri[t][2][1]);
          if (d < 0.f) { int tt = i1; i1 = i2; i2 = tt; }   // normal must point away from the inside corners
          faces[(out + n + t) * 3] = i0; faces[(out + n + t) * 3 + 1] = i1; faces[(out + n + t) * 3 + 2] = i2;
        }
      }
      n += ntri;
    }
  }
  if (!WRITE) tcount[c] = n;
}

// Marching cubes: the cell's 8 corner states select a row of the generated table; a triangle corner is the crossing on
// one of the 12 cube edges, i.e. the vertex of an axis edge of one of the cell's corners.
template <bool WRITE>
__global__ void k_mc_cells(const float* __restrict__ sdf, int D, float level,
                           const unsigned char* __restrict__ vmask, const int* __restrict__ vbase,
                           int* __restrict__ tcount, const int* __restrict__ tbase, int* __restrict__ faces) {
  int c = blockIdx.x * blockDim.x + threadIdx.x;
  const int C = D - 1;
  if (c >= C * C * C) return;
  int z = c % C, y = (c / C) % C, x = c / (C * C);
  unsigned inside = 0;
#pragma unroll
  for (int b = 0; b < 8; ++b)
    if (sdf[mt_lin(x + (b & 1), y + ((b >> 1) & 1), z + ((b >> 2) & 1), D)] < level) inside |= 1u << b;
  const int n = c_mc_ntri[inside];
  if (!WRITE) { tcount[c] = n; return; }
  int out = tbase[c];
  for (int t = 0; t < n; ++t)
#pragma unroll
    for (int k = 0; k < 3; ++k) {
      const int e = c_mc_tri[inside][3 * t + k], a = c_mc_edges[e][0], b = c_mc_edges[e][1];
      const int v = mt_lin(x + (a & 1), y + ((a >> 1) & 1), z + ((a >> 2) & 1), D);
      const int ty = (b ^ a) - 1;                          // axis edge type of corner a: 0, 1 or 3
      faces[(out + t) * 3 + k] = vbase[v] + __popc(vmask[v] & ((1u << ty) - 1u));
    }
}

// ------------------------------------------------------------------------------------------------
// exclusive scan of n ints over many blocks (the grids have millions of entries): block-local scan of
// 1024-entry segments + segment totals, scan of the totals (k_exclusive_scan, one block), add back.
// out[0..n] with out[n] = total; deterministic.
// ------------------------------------------------------------------------------------------------
#define SCAN_SEG 1024
__global__ void __launch_bounds__(SCAN_SEG) k_scan_segments(const int* __restrict__ in, int* __restrict__ out, int n,
                                                            int* __restrict__ seg_sum) {
  __shared__ int wsum[32];
  const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
  const long i = (long)blockIdx.x * SCAN_SEG + tid;
  const int v = i < n ? in[i] : 0;
  int x = v;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    int y = __shfl_up_sync(0xffffffffu, x, d);
    if (lane >= d) x += y;
  }
  if (lane == 31) wsum[w] = x;
  __syncthreads();
  if (w == 0) {
    int s = wsum[lane], t = s;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      int y = __shfl_up_sync(0xffffffffu, t, d);
      if (lane >= d) t += y;
    }
    wsum[lane] = t - s;
    if (lane == 31) seg_sum[blockIdx.x] = t;
  }
  __syncthreads();
  if (i < n) out[i] = wsum[w] + x - v;
}

__global__ void __launch_bounds__(SCAN_SEG) k_scan_add(int* __restrict__ out, int n, const int* __restrict__ seg_base,
                                                       int nseg) {
  const long i = (long)blockIdx.x * SCAN_SEG + threadIdx.x;
  if (i < n) out[i] += seg_base[blockIdx.x];
  if (i == 0) out[n] = seg_base[nseg];
}

// ------------------------------------------------------------------------------------------------
// Grid decode for the isosurface: only grid points that are corners of a cell the level set may cross need the
// fp32-equivalent SDF - the isosurface kernels interpolate on edges with a sign change and never look at the
// values of other cells.  A cell is ACTIVE when min(corner) - level < margin and max(corner) - level > -margin on
// the first-pass values (|first pass - exact| < margin): every edge with a sign change, in the first-pass or in
// the exact values, lies in active cells only, so the mesh of the mixed grid equals the mesh of the exact grid.
// ------------------------------------------------------------------------------------------------
__global__ void k_grid_mark(const float* __restrict__ sdf, int D, float level, float margin,
                            unsigned char* __restrict__ need) {
  const long c = (long)blockIdx.x * blockDim.x + threadIdx.x;
  const int E = D - 1;
  if (c >= (long)E * E * E) return;
  const int z = (int)(c % E), y = (int)((c / E) % E), x = (int)(c / ((long)E * E));
  const long b = ((long)x * D + y) * D + z;
  float lo = 1e30f, hi = -1e30f;
#pragma unroll
  for (int k = 0; k < 8; ++k) {
    const float v = sdf[b + ((k >> 2) & 1) * (long)D * D + ((k >> 1) & 1) * D + (k & 1)] - level;
    lo = fminf(lo, v); hi = fmaxf(hi, v);
  }
  if (lo < margin && hi > -margin) {
#pragma unroll
    for (int k = 0; k < 8; ++k) need[b + ((k >> 2) & 1) * (long)D * D + ((k >> 1) & 1) * D + (k & 1)] = 1;
  }
}

// row list of the marked grid points (order irrelevant: the decoder gathers and scatters through it)
__global__ void k_grid_list(const unsigned char* __restrict__ need, long V, int* __restrict__ list,
                            int* __restrict__ count) {
  const long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
  const bool f = i < V && need[i];
  const unsigned m = __ballot_sync(0xffffffffu, f);
  if (!m) return;
  const int lane = threadIdx.x & 31;
  int base = 0;
  if (lane == __ffs(m) - 1) base = atomicAdd(count, __popc(m));
  base = __shfl_sync(0xffffffffu, base, __ffs(m) - 1);
  if (f) list[base + __popc(m & ((1u << lane) - 1u))] = (int)i;
}
