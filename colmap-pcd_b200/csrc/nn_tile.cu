// nn_tile.cu -- tile-staged exact 1-NN (a4/a5: Kdtree::GetClosestPoint, src/lidar/kdtree.cc:10-21; float32 L2 in
// FLANN's L2_Simple order (dx2+dy2)+dz2 on the float-cast query, src/lidar/ply.cc:92; ties -> lowest original index).
//
// A TILE is one 4 x 4 x 4 block of cubic cells (the unit of the occupancy words).  Per batch:
//   k_tile_count   query -> tile histogram (one global REDG per query)
//   cub scan       tile_start
//   k_tile_jobs    tiles with >= tile_min queries are cut into jobs of <= TL_JOBQ queries
//   k_tile_scatter one-pass counting sort: every query takes the next slot of its tile (record + original index);
//                  queries of sparser tiles go straight to the far queue
//   k_nn_tile      persistent blocks draw jobs.  A block STAGES the tile and its one-cell halo in shared memory:
//                  the (y, z) rows of the region are contiguous runs of the cell-sorted points, so staging is one
//                  row of cell_start (coalesced) and one 1-D bulk copy (cp.async.bulk -> mbarrier) per row.  Every
//                  table lookup and every candidate load of the search then hits shared memory.
// Search of a job: each warp takes 32 queries at a time.  The scan is FLATTENED so that all 32 lanes stay busy:
// in three rounds (own cell; rest of the centre row inside the chord of the bound found; the eight ring rows inside
// their chords) every lane PLANS the candidate ranges of its own query from the local cell table, the ranges are cut
// into aligned 8-point blocks that are dealt round-robin to the lanes of the warp, and a lane scans whole blocks for
// whichever query they belong to, merging into that query's best with a shared-memory atomicMin.  Lane L reads point
// (c + L) & 7 of its block in step c, so the eight lanes of a quarter-warp always hit eight different 16-byte bank
// groups: the LDS.128 stream is conflict-free whatever the blocks are.  Aligned blocks may hold a few points outside
// the planned range; every staged point is a real cloud point, so scanning it can only help.
// Exactness is the same argument as in nn.cu: a range is skipped only when a conservative lower bound on its distance
// exceeds the current best; the winner is min over (d2 bits << 32 | original index); queries the K1 = 1 block cannot
// certify go to the far queue with their state.
#include "nn_dev.cuh"
#include <cub/cub.cuh>
#include <stdlib.h>

namespace pcd {

constexpr int TL_THREADS = 128;          // 4 warps share one staged tile
constexpr int TL_WARPS = TL_THREADS / 32;
constexpr int TL_JOBQ = 512;             // queries per job (a tile with more is cut into several jobs)
#ifndef TL_CAP
#define TL_CAP 2048                      // staged points per tile region; a region with more sends its queries to the far queue
#endif
constexpr int TL_K1 = 1;                 // halo in cubic cells
constexpr int TL_NY = 4 + 2 * TL_K1;     // rows of the region along y and z
constexpr int TL_ROWS = TL_NY * TL_NY;
constexpr int TL_MAXXS = 8;              // x sub-division the local table is sized for
constexpr int TL_MAXNX = TL_MAXXS * TL_NY;
constexpr int TL_MAXITEMS = 256;        // block items of one warp handled per pass of a round

struct TileJob { uint32_t tile, begin, count, pad; };

struct TileSmem {
  float4 pts[TL_CAP + 8];
  uint32_t tab[TL_ROWS * (TL_MAXNX + 1)];
  uint32_t row_base[TL_ROWS], row_off[TL_ROWS + 1];
  unsigned long long bar;
  uint32_t job, total;
  float4 rec[TL_WARPS][32];
  unsigned long long best[TL_WARPS][32];
  uint32_t items[TL_WARPS][TL_MAXITEMS];
};

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ uint32_t tile_of(const GridView& g, const float4& r) {
  const int ix = clampi((int)floorf((r.x - g.ox) * g.inv_cx), 0, g.dx - 1);
  const int iy = clampi((int)floorf((r.y - g.oy) * g.inv_c), 0, g.dy - 1);
  const int iz = clampi((int)floorf((r.z - g.oz) * g.inv_c), 0, g.dz - 1);
  return (uint32_t)((((iz >> 2) * g.nby) + (iy >> 2)) * g.nbx + ((ix / g.xs) >> 2));
}

__global__ void k_tile_count(const double* __restrict__ q, RadiusSrc rs, int64_t n, GridView g, uint32_t* __restrict__ cnt) {
  const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i >= n) return;
  const float4 r = query_record(q, rs, i);
  atomicAdd(&cnt[tile_of(g, r)], 1u);
}

// jobs of the tiles that are searched from shared memory; ctr[0] = number of jobs
__global__ void k_tile_jobs(const uint32_t* __restrict__ cnt, const uint32_t* __restrict__ start, int64_t n_tiles, uint32_t tile_min,
                            TileJob* __restrict__ jobs, unsigned int* __restrict__ ctr) {
  const int64_t t = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (t >= n_tiles) return;
  const uint32_t c = cnt[t];
  if (c < tile_min || c == 0) return;
  const uint32_t nj = (c + TL_JOBQ - 1) / TL_JOBQ, per = (c + nj - 1) / nj;
  uint32_t base = atomicAdd(&ctr[0], nj);
  uint32_t b = start[t];
  for (uint32_t k = 0; k < nj; k++) {
    const uint32_t m = min(per, c - k * per);
    jobs[base + k] = TileJob{(uint32_t)t, b, m, 0u};
    b += m;
  }
}

// `cursor` starts as a copy of tile_start; a query of a tile below tile_min is queued for the far search as is
__global__ void k_tile_scatter(const double* __restrict__ q, RadiusSrc rs, int64_t n, GridView g, const uint32_t* __restrict__ cnt,
                               uint32_t* __restrict__ cursor, uint32_t tile_min, float4* __restrict__ srec, uint32_t* __restrict__ sid,
                               FarItem* __restrict__ far, unsigned int* __restrict__ far_count) {
  const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  bool sparse = false;
  float4 r = make_float4(0, 0, 0, 0);
  if (i < n) {
    r = query_record(q, rs, i);
    const uint32_t t = tile_of(g, r);
    if (cnt[t] < tile_min) sparse = true;
    else {
      const uint32_t pos = atomicAdd(&cursor[t], 1u);
      srec[pos] = r;
      sid[pos] = (uint32_t)i;
    }
  }
  const unsigned m = __ballot_sync(0xffffffffu, sparse);
  if (m) {
    const int lane = threadIdx.x & 31, leader = __ffs(m) - 1;
    unsigned base = 0;
    if (lane == leader) base = atomicAdd(far_count, (unsigned)__popc(m));
    base = __shfl_sync(0xffffffffu, base, leader);
    if (sparse) {
      FarItem it; it.best = ~0ull; it.i = (uint32_t)i; it.pad = 0; it.rec = r;
      far[base + __popc(m & ((1u << lane) - 1u))] = it;
    }
  }
}

// ---- mbarrier / bulk-copy wrappers (PTX ISA: mbarrier, cp.async.bulk) ----
__device__ __forceinline__ void mbar_init(unsigned long long* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(unsigned long long* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned long long* bar, uint32_t parity) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "WAIT_%=:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra DONE_%=;\n"
      "bra WAIT_%=;\n"
      "DONE_%=:\n"
      "}\n" ::"r"(smem_u32(bar)), "r"(parity) : "memory");
}
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, uint32_t bytes, unsigned long long* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst)), "l"(src),
               "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}
__device__ __forceinline__ void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

// One aligned block of eight staged points against one query; lane-rotated so that a quarter-warp never collides.
__device__ __forceinline__ unsigned long long scan_block8(const float4* __restrict__ blk, int rot, float qx, float qy, float qz,
                                                          unsigned long long bst) {
#pragma unroll
  for (int c = 0; c < 8; c++) {
    const float4 p = blk[(c + rot) & 7];
    const float ddx = qx - p.x, ddy = qy - p.y, ddz = qz - p.z;
    const float d = __fadd_rn(__fadd_rn(__fmul_rn(ddx, ddx), __fmul_rn(ddy, ddy)), __fmul_rn(ddz, ddz));
    const unsigned long long key = ((unsigned long long)__float_as_uint(d) << 32) | __float_as_uint(p.w);
    bst = key < bst ? key : bst;
  }
  return bst;
}

// Deals the planned block ranges of the warp's 32 queries to its lanes and scans them.  Each lane passes up to NR
// ranges [bs[k], be[k]) of 8-point blocks for ITS query (slot = lane); returns after every item has been merged.
template <int NR>
__device__ __forceinline__ void flat_round(TileSmem& S, int w, int lane, const uint32_t (&bs)[NR], const uint32_t (&be)[NR]) {
  uint32_t mine = 0;
#pragma unroll
  for (int k = 0; k < NR; k++) mine += be[k] - bs[k];
  uint32_t incl = mine;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const uint32_t v = __shfl_up_sync(0xffffffffu, incl, o);
    if (lane >= o) incl += v;
  }
  const uint32_t total = __shfl_sync(0xffffffffu, incl, 31);
  if (total == 0) return;
  uint32_t* items = S.items[w];
  for (uint32_t done = 0; done < total; done += TL_MAXITEMS) {      // more than TL_MAXITEMS blocks: several passes
    uint32_t pos = incl - mine;                                      // global item number of this lane's first block
#pragma unroll
    for (int k = 0; k < NR; k++) {
      for (uint32_t b = bs[k]; b < be[k]; b++, pos++)
        if (pos >= done && pos < done + TL_MAXITEMS) items[pos - done] = ((uint32_t)lane << 16) | b;
    }
    __syncwarp();
    const uint32_t cnt = min(total - done, (uint32_t)TL_MAXITEMS);
    for (uint32_t it = lane; it < cnt; it += 32) {
      const uint32_t item = items[it];
      const int slot = (int)(item >> 16);
      const float4 r = S.rec[w][slot];
      const unsigned long long cur = S.best[w][slot];
      const unsigned long long bst = scan_block8(&S.pts[(item & 0xffffu) << 3], lane & 7, r.x, r.y, r.z, cur);
      if (bst < cur) atomicMin(&S.best[w][slot], bst);
    }
    __syncwarp();
  }
}

__global__ void __launch_bounds__(TL_THREADS) k_nn_tile(GridView g, const TileJob* __restrict__ jobs, const unsigned int* __restrict__ ctr_in,
                                                        unsigned int* __restrict__ job_next, const float4* __restrict__ srec,
                                                        const uint32_t* __restrict__ sid, int32_t* __restrict__ idx_out,
                                                        float* __restrict__ d2_out, FarItem* __restrict__ far,
                                                        unsigned int* __restrict__ far_count) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  TileSmem& S = *reinterpret_cast<TileSmem*>(smem_raw);
  const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
  const uint32_t njobs = ctr_in[0];
  const int NX = g.xs * TL_NY, NXE = NX + 1;      // x cells of the region, table entries per row
  const int kx = TL_K1 * g.xs;
  if (tid == 0) mbar_init(&S.bar, 1);
  fence_async_smem();
  __syncthreads();
  uint32_t parity = 0;
  for (;;) {
    if (tid == 0) S.job = atomicAdd(job_next, 1u);
    __syncthreads();                                // also: every read of the previous region is done
    const uint32_t j = S.job;
    if (j >= njobs) return;
    const TileJob job = jobs[j];
    const int X = (int)(job.tile % (uint32_t)g.nbx), Y = (int)((job.tile / (uint32_t)g.nbx) % (uint32_t)g.nby),
              Z = (int)(job.tile / ((uint32_t)g.nbx * (uint32_t)g.nby));
    const int x0r = X * 4 * g.xs - kx, y0r = 4 * Y - TL_K1, z0r = 4 * Z - TL_K1;
    // ---- stage 1: the region's rows of cell_start (global offsets) ----
    for (int r = w; r < TL_ROWS; r += TL_WARPS) {
      const int y = y0r + r % TL_NY, z = z0r + r / TL_NY;
      const bool in = y >= 0 && y < g.dy && z >= 0 && z < g.dz;
      const size_t lin = ((size_t)(in ? z : 0) * g.dy + (in ? y : 0)) * g.dx;
      for (int k = lane; k < NXE; k += 32) {
        const int x = clampi(x0r + k, 0, g.dx);
        S.tab[r * NXE + k] = in ? __ldg(&g.cell_start[lin + x]) : 0u;
      }
    }
    __syncthreads();
    if (w == 0) {                                    // row lengths -> offsets of the rows in the staged array
      uint32_t len0 = 0, len1 = 0, base0 = 0, base1 = 0;
      if (lane < TL_ROWS) { base0 = S.tab[lane * NXE]; len0 = S.tab[lane * NXE + NX] - base0; }
      if (lane + 32 < TL_ROWS) { base1 = S.tab[(lane + 32) * NXE]; len1 = S.tab[(lane + 32) * NXE + NX] - base1; }
      uint32_t i0 = len0, i1 = len1;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const uint32_t a = __shfl_up_sync(0xffffffffu, i0, o), b = __shfl_up_sync(0xffffffffu, i1, o);
        if (lane >= o) { i0 += a; i1 += b; }
      }
      const uint32_t t0 = __shfl_sync(0xffffffffu, i0, 31);
      i1 += t0;
      const uint32_t total = __shfl_sync(0xffffffffu, i1, 31);
      if (lane < TL_ROWS) { S.row_base[lane] = base0; S.row_off[lane] = i0 - len0; }
      if (lane + 32 < TL_ROWS) { S.row_base[lane + 32] = base1; S.row_off[lane + 32] = i1 - len1; }
      if (lane == 0) {
        S.row_off[TL_ROWS] = total;
        S.total = total;
        if (total > 0 && total <= TL_CAP) mbar_expect_tx(&S.bar, total * 16u);
      }
    }
    __syncthreads();
    const uint32_t total = S.total;
    const bool staged = total <= TL_CAP;
    if (staged) {
      // ---- stage 2: one bulk copy per non-empty row; meanwhile the table becomes local offsets ----
      if (tid < TL_ROWS) {
        const uint32_t len = S.row_off[tid + 1] - S.row_off[tid];
        if (len) { fence_async_smem(); bulk_g2s(&S.pts[S.row_off[tid]], g.pts + S.row_base[tid], len * 16u, &S.bar); }
      }
      for (int e = tid; e < TL_ROWS * NXE; e += TL_THREADS) {
        const int r = e / NXE;
        S.tab[e] = S.tab[e] - S.row_base[r] + S.row_off[r];
      }
      if (total) { mbar_wait(&S.bar, parity); parity ^= 1u; }
      // the last block is completed with copies of the last point (duplicates cannot change a minimum)
      if (total && tid < 8 && total + tid < ((total + 7u) & ~7u)) S.pts[total + tid] = S.pts[total - 1];
      __syncthreads();
    }
    // ---- search: each warp takes 32 queries of the job at a time ----
    const uint32_t chunks = (job.count + 31u) / 32u, per = (job.count + chunks - 1u) / chunks;
    for (uint32_t ch = w; ch < chunks; ch += TL_WARPS) {
      const uint32_t q0 = ch * per, qn = min(per, job.count - q0);
      const bool has = (uint32_t)lane < qn;
      float4 rec = make_float4(0.f, 0.f, 0.f, 0.f);
      uint32_t orig = 0;
      if (has) { rec = __ldg(&srec[job.begin + q0 + lane]); orig = __ldg(&sid[job.begin + q0 + lane]); }
      Query Q;
      const bool valid = has && Q.from_record(g, rec);
      bool resolved = true;
      if (!staged) resolved = !valid;                // region too dense for shared memory: far search from scratch
      else {
        S.rec[w][lane] = rec;
        S.best[w][lane] = ~0ull;
        __syncwarp();
        const int lx = Q.ix - x0r, ly = Q.iy - y0r, lz = Q.iz - z0r;
        const uint32_t* row0 = &S.tab[(lz * TL_NY + ly) * NXE];
        // round 0: own cell
        uint32_t b0 = 0, e0 = 0;
        if (valid) { b0 = row0[lx]; e0 = row0[lx + 1]; }
        uint32_t B0s = b0 >> 3, B0e = e0 > b0 ? (e0 + 7) >> 3 : B0s;
        {
          const uint32_t bs[1] = {B0s}, be[1] = {B0e};
          flat_round<1>(S, w, lane, bs, be);
        }
        // round 1: rest of the centre row inside the chord of the bound
        {
          uint32_t bs[2] = {0, 0}, be[2] = {0, 0};
          if (valid) {
            Q.best = S.best[w][lane];
            int x0 = Q.ix - kx, x1 = Q.ix + kx;
            Q.chord(g, Q.bound2(), 0.0f, x0, x1);
            x0 = min(x0, Q.ix); x1 = max(x1, Q.ix);
            const uint32_t b = row0[x0 - x0r], e = row0[x1 + 1 - x0r];
            const uint32_t Bs = b >> 3, Be = e > b ? (e + 7) >> 3 : Bs;
            // blocks of [b, e) not yet scanned: those below B0s and those from B0e on
            if (e0 > b0) { bs[0] = Bs; be[0] = max(Bs, min(B0s, Be)); bs[1] = min(max(B0e, Bs), Be); be[1] = Be; }
            else { bs[0] = Bs; be[0] = Be; }
          }
          flat_round<2>(S, w, lane, bs, be);
        }
        // round 2: the eight ring rows inside their chords
        {
          uint32_t bs[8], be[8];
#pragma unroll
          for (int k = 0; k < 8; k++) { bs[k] = 0; be[k] = 0; }
          if (valid) {
            Q.best = S.best[w][lane];
            const float b2 = Q.bound2();
            float gy2[3], gz2[3];
#pragma unroll
            for (int d = -1; d <= 1; d++) {
              const float a = Query::gap(Q.py, Q.iy + d, Q.iy + d, g.c, g.margin), b = Query::gap(Q.pz, Q.iz + d, Q.iz + d, g.c, g.margin);
              gy2[d + 1] = (Q.iy + d < 0 || Q.iy + d >= g.dy) ? __int_as_float(0x7fc00000) : a * a;
              gz2[d + 1] = (Q.iz + d < 0 || Q.iz + d >= g.dz) ? __int_as_float(0x7fc00000) : b * b;
            }
            int k = 0;
#pragma unroll
            for (int dz = -1; dz <= 1; dz++)
#pragma unroll
              for (int dy = -1; dy <= 1; dy++) {
                if (dy == 0 && dz == 0) continue;
                const float g2 = (gy2[dy + 1] + gz2[dz + 1]) * SHRINK;
                if (g2 <= b2) {
                  int x0 = Q.ix - kx, x1 = Q.ix + kx;
                  if (Q.chord(g, b2, g2, x0, x1)) {
                    const uint32_t* row = &S.tab[((lz + dz) * TL_NY + (ly + dy)) * NXE];
                    const uint32_t b = row[x0 - x0r], e = row[x1 + 1 - x0r];
                    if (e > b) { bs[k] = b >> 3; be[k] = (e + 7) >> 3; }
                  }
                }
                k++;
              }
          }
          flat_round<8>(S, w, lane, bs, be);
        }
        if (valid) {
          Q.best = S.best[w][lane];
          const float rc2 = Q.certified2(g, kx, TL_K1);
          resolved = Q.best_d2() <= rc2 || Q.R2 <= rc2;
        }
        __syncwarp();
      }
      if (has && resolved) nn_write(Q.best, orig, idx_out, d2_out);
      const unsigned m = __ballot_sync(0xffffffffu, has && !resolved);
      if (m) {
        const int leader = __ffs(m) - 1;
        unsigned base = 0;
        if (lane == leader) base = atomicAdd(far_count, (unsigned)__popc(m));
        base = __shfl_sync(0xffffffffu, base, leader);
        if (has && !resolved) {
          FarItem it; it.best = Q.best; it.i = orig; it.pad = staged ? 1u : 0u; it.rec = Q.record();
          far[base + __popc(m & ((1u << lane) - 1u))] = it;
        }
      }
    }
  }
}

// Host side of the tile path: fills the far queue with what the tiles could not resolve.  Returns PCD_OK or an error;
// *used = false when the grid does not suit the tile path (caller falls back to the cell-sorted search).
int nn_tile_search(pcd_cloud* cloud, const double* d_q, const RadiusSrc& rs, int64_t n, int32_t* d_idx, float* d2,
                   FarItem* far, unsigned int* far_count, cudaStream_t st, bool* used) {
  pcd_ctx* ctx = cloud->ctx;
  const GridView& g = cloud->grid;
  *used = false;
  if (g.xs > TL_MAXXS) return PCD_OK;
  const int64_t n_tiles = (int64_t)g.nbx * g.nby * g.nbz;
  if (n_tiles >= ((int64_t)1 << 31)) return PCD_OK;
  *used = true;
  uint32_t tile_min = 8;
  if (const char* e = getenv("PCD_B200_TILE_MIN")) tile_min = (uint32_t)atoi(e);
  if (tile_min < 1) tile_min = 1;
  static bool attr_set = false;
  if (!attr_set) {
    PCD_CUDA(ctx, cudaFuncSetAttribute(k_nn_tile, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(TileSmem)));
    attr_set = true;
  }
  static int occ = 0;
  if (occ == 0) {
    PCD_CUDA(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, k_nn_tile, TL_THREADS, sizeof(TileSmem)));
    if (occ < 1) occ = 1;
  }
  const size_t max_jobs = (size_t)n / tile_min + (size_t)n / TL_JOBQ + 2;
  PCD_CUDA(ctx, ctx->tile_cnt.ensure(4 * (size_t)(n_tiles + 1)));
  PCD_CUDA(ctx, ctx->tile_start.ensure(4 * (size_t)(n_tiles + 1)));
  PCD_CUDA(ctx, ctx->tile_jobs.ensure(sizeof(TileJob) * max_jobs + 16));
  PCD_CUDA(ctx, ctx->srec.ensure(16 * (size_t)n));
  PCD_CUDA(ctx, ctx->sid.ensure(4 * (size_t)n));
  uint32_t* cnt = ctx->tile_cnt.as<uint32_t>();
  uint32_t* start = ctx->tile_start.as<uint32_t>();
  unsigned int* ctr = ctx->tile_jobs.as<unsigned int>();              // [0] jobs, [1] next job
  TileJob* jobs = (TileJob*)(ctx->tile_jobs.as<char>() + 16);
  {
    ProfScope ps(ctx, PCD_K_NN_KEYSORT, st);
    PCD_CUDA(ctx, cudaMemsetAsync(cnt, 0, 4 * (size_t)(n_tiles + 1), st));
    PCD_CUDA(ctx, cudaMemsetAsync(ctr, 0, 16, st));
    k_tile_count<<<pcd_blocks(n, 256), 256, 0, st>>>(d_q, rs, n, g, cnt);
    PCD_LAUNCH_CHECK(ctx);
    size_t tmp = 0;
    PCD_CUDA(ctx, cub::DeviceScan::ExclusiveSum(nullptr, tmp, cnt, start, (int)(n_tiles + 1), st));
    PCD_CUDA(ctx, ctx->cub_tmp.ensure(tmp));
    PCD_CUDA(ctx, cub::DeviceScan::ExclusiveSum(ctx->cub_tmp.p, tmp, cnt, start, (int)(n_tiles + 1), st));
    k_tile_jobs<<<pcd_blocks(n_tiles, 256), 256, 0, st>>>(cnt, start, n_tiles, tile_min, jobs, ctr);
    PCD_LAUNCH_CHECK(ctx);
    k_tile_scatter<<<pcd_blocks(n, 256), 256, 0, st>>>(d_q, rs, n, g, cnt, start, tile_min, ctx->srec.as<float4>(), ctx->sid.as<uint32_t>(),
                                                       far, far_count);
    PCD_LAUNCH_CHECK(ctx);
  }
  {
    ProfScope ps(ctx, PCD_K_NN_SEARCH, st);
    int blocks = ctx->sm_count * occ;
    k_nn_tile<<<blocks, TL_THREADS, sizeof(TileSmem), st>>>(g, jobs, ctr, ctr + 1, ctx->srec.as<float4>(), ctx->sid.as<uint32_t>(), d_idx, d2,
                                                            far, far_count);
    PCD_LAUNCH_CHECK(ctx);
  }
  return PCD_OK;
}

}  // namespace pcd
