// CPU simulator of grid-search strategies (round 2 design study): counts candidates per query and models the warp cost of
// the row loops (max over lanes per scan site), a per-lane state machine and a perfectly flattened scan.  Input: pts.f32, q.f32,
// rad.f32 written by tools/nn_sim_gen.py.  usage: nn_sim cell xs K1 variant SW skipbits (skipbits < 0: tile-sorted order)
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <math.h>
#include <stdint.h>

typedef struct { float x, y, z; uint32_t id; } P4;
static float ox, oy, oz, c, inv_c, cx, inv_cx, margin;
static int xs, dx, dy, dz;
static uint32_t* cell_start; static P4* pts; static int64_t ncells;
static const float SHRINK = 0.999999f, GROW = 1.000001f;

static void* rd(const char* fn, size_t* n) { FILE* f = fopen(fn, "rb"); fseek(f, 0, SEEK_END); size_t s = ftell(f); fseek(f, 0, SEEK_SET); void* p = malloc(s); if (fread(p, 1, s, f) != s) exit(1); fclose(f); *n = s; return p; }
static int clampi(int v, int lo, int hi) { return v < lo ? lo : (v > hi ? hi : v); }

typedef struct {
  float qx, qy, qz, px, py, pz, R2; int ix, iy, iz; float bd2; uint32_t bi;
  int nsite; int len[64];   // candidates scanned per site (0 = pruned or empty)
  int visited;              // sites that reached the cell_start loads
} Qs;

static void scan(Qs* q, uint32_t b, uint32_t e) {
  for (uint32_t j = b; j < e; j++) {
    float ddx = q->qx - pts[j].x, ddy = q->qy - pts[j].y, ddz = q->qz - pts[j].z;
    float d = (ddx * ddx + ddy * ddy) + ddz * ddz;
    if (d < q->bd2 || (d == q->bd2 && pts[j].id < q->bi)) { q->bd2 = d; q->bi = pts[j].id; }
  }
}
static float gapf(float p, int lo, int hi, float cc) { float a = fmaxf((float)lo * cc - p, p - (float)(hi + 1) * cc); return fmaxf(a - margin, 0.0f); }
static float bound2(Qs* q) { return fminf(q->bd2, q->R2); }
static int chord(Qs* q, float b2, float g2, int* x0, int* x1) {
  float hx = sqrtf(b2 - g2) * GROW + margin;
  int a = (int)floorf((q->px - hx) * inv_cx), b = (int)floorf((q->px + hx) * inv_cx);
  if (*x0 < 0) *x0 = 0; if (a > *x0) *x0 = a;
  if (*x1 > dx - 1) *x1 = dx - 1; if (b < *x1) *x1 = b;
  return *x0 <= *x1;
}
static void site(Qs* q, uint32_t b, uint32_t e) { q->len[q->nsite++] = (int)(e - b); q->visited++; scan(q, b, e); }
static void skip(Qs* q) { q->len[q->nsite++] = 0; }

// excl: x cells [ex0, ex1] of this row were already scanned (seed stage); pass ex0 > ex1 for none
static void row(Qs* q, int kx, int ddy, int ddz, int ex0, int ex1) {
  int y = q->iy + ddy, z = q->iz + ddz;
  if (y < 0 || y >= dy || z < 0 || z >= dz) { skip(q); return; }
  float a = gapf(q->py, y, y, c), b = gapf(q->pz, z, z, c);
  float g2 = (a * a + b * b) * SHRINK, b2 = bound2(q);
  if (!(g2 <= b2)) { skip(q); return; }
  int x0 = q->ix - kx, x1 = q->ix + kx;
  if (!chord(q, b2, g2, &x0, &x1)) { skip(q); return; }
  size_t lin = ((size_t)z * dy + y) * dx;
  if (ex0 <= ex1) {
    uint32_t n = 0; q->visited++;
    if (x0 < ex0) { uint32_t bb = cell_start[lin + x0], ee = cell_start[lin + ex0]; n += ee - bb; scan(q, bb, ee); }
    if (x1 > ex1) { uint32_t bb = cell_start[lin + ex1 + 1], ee = cell_start[lin + x1 + 1]; n += ee - bb; scan(q, bb, ee); }
    q->len[q->nsite++] = (int)n;
  } else site(q, cell_start[lin + x0], cell_start[lin + x1 + 1]);
}
static float certified2(Qs* q, int kx, int k) {
  float rc = INFINITY;
  if (q->ix - kx > 0) rc = fminf(rc, q->px - (float)(q->ix - kx) * cx);
  if (q->ix + kx < dx - 1) rc = fminf(rc, (float)(q->ix + kx + 1) * cx - q->px);
  if (q->iy - k > 0) rc = fminf(rc, q->py - (float)(q->iy - k) * c);
  if (q->iy + k < dy - 1) rc = fminf(rc, (float)(q->iy + k + 1) * c - q->py);
  if (q->iz - k > 0) rc = fminf(rc, q->pz - (float)(q->iz - k) * c);
  if (q->iz + k < dz - 1) rc = fminf(rc, (float)(q->iz + k + 1) * c - q->pz);
  if (rc == INFINITY) return INFINITY;
  rc = fmaxf(rc - margin, 0.0f);
  return rc * rc * SHRINK;
}

// ring enumeration like the kernel: for k: t=0..k: 8 sign combos
static int ring_rows(int k, int out[][2]) {
  int n = 0;
  for (int t = 0; t <= k; t++)
    for (int s = 0; s < 8; s++) {
      int a = (s & 1) ? k : -k, bs = (s & 2) ? t : -t;
      int ddy = (s & 4) ? a : bs, ddz = (s & 4) ? bs : a;
      if (t == 0 && (s & 2)) continue;
      if (t == k && (s & 4)) continue;
      out[n][0] = ddy; out[n][1] = ddz; n++;
    }
  return n;
}

// variant 0: current kernel.  variant 1: seed = own x cell +- SW in all rows of ring<=KS (nearest first, pruned), then chords.
static int search(Qs* q, int K1, int variant, int SW) {
  const int kx = K1 * xs;
  q->nsite = 0; q->visited = 0; q->bd2 = INFINITY; q->bi = 0xffffffffu;
  size_t lin0 = ((size_t)q->iz * dy + q->iy) * dx;
  int rr[64][2];
  if (variant == 0) {
    uint32_t b0 = cell_start[lin0 + q->ix], e0 = cell_start[lin0 + q->ix + 1];
    site(q, b0, e0);
    int x0 = q->ix - kx, x1 = q->ix + kx;
    chord(q, bound2(q), 0.0f, &x0, &x1);
    if (x0 > q->ix) x0 = q->ix; if (x1 < q->ix) x1 = q->ix;
    uint32_t b = cell_start[lin0 + x0], e = cell_start[lin0 + x1 + 1];
    site(q, b, b0); site(q, e0, e);
    for (int k = 1; k <= K1; k++) {
      int n = ring_rows(k, rr);
      for (int i = 0; i < n; i++) row(q, kx, rr[i][0], rr[i][1], 1, 0);
      float rc2 = certified2(q, kx, k);
      if (q->bd2 <= rc2 || q->R2 <= rc2) return 1;
    }
    return 0;
  }
  if (variant == 4) {
    // own cell; axis rows' own x cell (pruned by gap <= bound); centre remainder; ring remainders progressive
    uint32_t b0 = cell_start[lin0 + q->ix], e0 = cell_start[lin0 + q->ix + 1];
    site(q, b0, e0);
    int ex0 = q->ix - SW < 0 ? 0 : q->ix - SW, ex1 = q->ix + SW > dx - 1 ? dx - 1 : q->ix + SW;
    int ax[4][2] = {{-1,0},{1,0},{0,-1},{0,1}};
    int seeded[3][3] = {{0}};
    for (int i = 0; i < 4; i++) {
      int y = q->iy + ax[i][0], z = q->iz + ax[i][1];
      if (y < 0 || y >= dy || z < 0 || z >= dz) { skip(q); continue; }
      float a = gapf(q->py, y, y, c), b = gapf(q->pz, z, z, c);
      float g2 = (a * a + b * b) * SHRINK;
      if (!(g2 <= bound2(q))) { skip(q); continue; }
      size_t lin = ((size_t)z * dy + y) * dx;
      site(q, cell_start[lin + ex0], cell_start[lin + ex1 + 1]); seeded[ax[i][0] + 1][ax[i][1] + 1] = 1;
    }
    { int x0 = q->ix - kx, x1 = q->ix + kx; chord(q, bound2(q), 0.0f, &x0, &x1); if (x0 > q->ix) x0 = q->ix; if (x1 < q->ix) x1 = q->ix;
      uint32_t b = cell_start[lin0 + x0], e = cell_start[lin0 + x1 + 1]; q->visited++; scan(q, b, b0); scan(q, e0, e); q->len[q->nsite++] = (int)((b0 - b) + (e - e0)); }
    int n = ring_rows(1, rr);
    for (int i = 0; i < n; i++) { if (seeded[rr[i][0] + 1][rr[i][1] + 1]) row(q, kx, rr[i][0], rr[i][1], ex0, ex1); else row(q, kx, rr[i][0], rr[i][1], 1, 0); }
    float rc2 = certified2(q, kx, 1);
    return (q->bd2 <= rc2 || q->R2 <= rc2);
  }
  if (variant == 3) {
    if (SW == 1) { uint32_t b0 = cell_start[lin0 + q->ix], e0 = cell_start[lin0 + q->ix + 1]; site(q, b0, e0); }
    for (int k = 0; k <= K1; k++) {
      int n = 1; rr[0][0] = 0; rr[0][1] = 0;
      if (k > 0) n = ring_rows(k, rr);
      float b2 = bound2(q);
      for (int i = 0; i < n; i++) {
        int y = q->iy + rr[i][0], z = q->iz + rr[i][1];
        if (y < 0 || y >= dy || z < 0 || z >= dz) continue;
        float a = gapf(q->py, y, y, c), b = gapf(q->pz, z, z, c);
        float g2 = (a * a + b * b) * SHRINK;
        if (!(g2 <= b2)) continue;
        int x0 = q->ix - kx, x1 = q->ix + kx;
        if (!chord(q, b2, g2, &x0, &x1)) continue;
        size_t lin = ((size_t)z * dy + y) * dx;
        uint32_t bb = cell_start[lin + x0], ee = cell_start[lin + x1 + 1];
        if (k == 0 && SW == 1) { uint32_t b0 = cell_start[lin0 + q->ix], e0 = cell_start[lin0 + q->ix + 1]; q->visited++; scan(q, bb, b0); scan(q, e0, ee); q->len[q->nsite++] = (int)(ee - bb - (e0 - b0)); }
        else site(q, bb, ee);
      }
      if (k > 0) { float rc2 = certified2(q, kx, k); if (q->bd2 <= rc2 || q->R2 <= rc2) return 1; }
    }
    return 0;
  }
  if (variant == 2) {
    int ex0 = clampi(q->ix - SW, 0, dx - 1), ex1 = clampi(q->ix + SW, 0, dx - 1);
    float b2 = q->R2; int nr = 0;
    for (int k = 0; k <= K1; k++) {
      int n = 1; rr[0][0] = 0; rr[0][1] = 0;
      if (k > 0) n = ring_rows(k, rr);
      for (int i = 0; i < n; i++) {
        int y = q->iy + rr[i][0], z = q->iz + rr[i][1];
        if (y < 0 || y >= dy || z < 0 || z >= dz) continue;
        float a = gapf(q->py, y, y, c), b = gapf(q->pz, z, z, c);
        float g2 = (a * a + b * b) * SHRINK;
        if (!(g2 <= b2)) continue;
        size_t lin = ((size_t)z * dy + y) * dx;
        uint32_t bb = cell_start[lin + ex0], ee = cell_start[lin + ex1 + 1];
        if (ee > bb) { q->len[0] += 0; nr++; }
        q->visited++; scan(q, bb, ee); q->len[q->nsite++] = (int)(ee - bb);
      }
    }
    b2 = bound2(q);
    for (int k = 0; k <= K1; k++) {
      int n = 1; rr[0][0] = 0; rr[0][1] = 0;
      if (k > 0) n = ring_rows(k, rr);
      for (int i = 0; i < n; i++) {
        int y = q->iy + rr[i][0], z = q->iz + rr[i][1];
        if (y < 0 || y >= dy || z < 0 || z >= dz) continue;
        float a = gapf(q->py, y, y, c), b = gapf(q->pz, z, z, c);
        float g2 = (a * a + b * b) * SHRINK;
        if (!(g2 <= b2)) continue;
        int x0 = q->ix - kx, x1 = q->ix + kx;
        if (!chord(q, b2, g2, &x0, &x1)) continue;
        size_t lin = ((size_t)z * dy + y) * dx;
        uint32_t nn = 0;
        if (x0 < ex0) { uint32_t bb = cell_start[lin + x0], ee = cell_start[lin + ex0]; nn += ee - bb; scan(q, bb, ee); }
        if (x1 > ex1) { uint32_t bb = cell_start[lin + ex1 + 1], ee = cell_start[lin + x1 + 1]; nn += ee - bb; scan(q, bb, ee); }
        if (x0 < ex0 || x1 > ex1) { q->visited++; q->len[q->nsite++] = (int)nn; }
      }
    }
    float rc2 = certified2(q, kx, K1);
    return (q->bd2 <= rc2 || q->R2 <= rc2);
  }
  // variant 1: seed stage over x cells [ix-SW, ix+SW] of every row in rings 0..K1 (nearest first), then remainder
  int ex0 = clampi(q->ix - SW, 0, dx - 1), ex1 = clampi(q->ix + SW, 0, dx - 1);
  for (int k = 0; k <= K1; k++) {
    int n = 1; rr[0][0] = 0; rr[0][1] = 0;
    if (k > 0) n = ring_rows(k, rr);
    for (int i = 0; i < n; i++) {
      int y = q->iy + rr[i][0], z = q->iz + rr[i][1];
      if (y < 0 || y >= dy || z < 0 || z >= dz) { skip(q); continue; }
      float a = gapf(q->py, y, y, c), b = gapf(q->pz, z, z, c);
      float g2 = (a * a + b * b) * SHRINK;
      if (!(g2 <= bound2(q))) { skip(q); continue; }
      size_t lin = ((size_t)z * dy + y) * dx;
      site(q, cell_start[lin + ex0], cell_start[lin + ex1 + 1]);
    }
  }
  for (int k = 0; k <= K1; k++) {
    int n = 1; rr[0][0] = 0; rr[0][1] = 0;
    if (k > 0) n = ring_rows(k, rr);
    for (int i = 0; i < n; i++) row(q, kx, rr[i][0], rr[i][1], ex0, ex1);
    if (k > 0) {
      float rc2 = certified2(q, kx, k);
      if (q->bd2 <= rc2 || q->R2 <= rc2) return 1;
    }
  }
  return 0;
}


static uint64_t* occw; static int nbx, nby, nbz;
static double cost_grp, cost_grpA, cost_grpB, cost_grpC;
static double f_blocks, f_words, f_nzwords, f_bits, f_cells, f_cands, f_n, f_shells, f_nobest;
static void far_block(Qs* q, float* b2, int X, int Y, int Z, int seen) {
  int sx = 4 * xs;
  f_blocks++;
  float a = gapf(q->px, X * sx, X * sx + sx - 1, cx), b = gapf(q->py, 4 * Y, 4 * Y + 3, c), cc = gapf(q->pz, 4 * Z, 4 * Z + 3, c);
  if (!((a * a + (b * b + cc * cc)) * SHRINK <= *b2)) return;
  uint64_t w = occw[((size_t)Z * nby + Y) * nbx + X]; f_words++;
  if (!w) return;
  f_nzwords++;
  for (int bit = 0; bit < 64; bit++) if ((w >> bit) & 1) {
    int lx = bit & 3, ly = (bit >> 2) & 3, lz = bit >> 4;
    int cxc = 4 * X + lx, cy = 4 * Y + ly, cz = 4 * Z + lz;
    float ga = gapf(q->px, cxc * xs, cxc * xs + xs - 1, cx), gb = gapf(q->py, cy, cy, c), gc = gapf(q->pz, cz, cz, c);
    if (!(ga * ga * SHRINK <= *b2 && gb * gb * SHRINK <= *b2 && gc * gc * SHRINK <= *b2)) continue;
    f_bits++;
    float yz2 = gb * gb + gc * gc;
    if (!((ga * ga + yz2) * SHRINK <= *b2)) continue;
    int x0 = cxc * xs, x1 = x0 + xs - 1;
    if (!chord(q, *b2, yz2 * SHRINK, &x0, &x1)) continue;
    size_t lin = ((size_t)cz * dy + cy) * dx;
    int pieces[2][2] = {{x0, x1}, {1, 0}};
    if (seen && abs(cy - q->iy) <= 1 && abs(cz - q->iz) <= 1) { pieces[0][1] = x1 < q->ix - xs - 1 ? x1 : q->ix - xs - 1; pieces[1][0] = x0 > q->ix + xs + 1 ? x0 : q->ix + xs + 1; pieces[1][1] = x1; }
    for (int p = 0; p < 2; p++) if (pieces[p][0] <= pieces[p][1]) {
      uint32_t bb = cell_start[lin + pieces[p][0]], ee = cell_start[lin + pieces[p][1] + 1];
      f_cells++; f_cands += ee - bb; scan(q, bb, ee); *b2 = bound2(q);
    }
  }
}
static float far_cert(Qs* q, int bx, int by, int bz, int kb) {
  int sx = 4 * xs; float rc = INFINITY;
  int x0 = (bx - kb) * sx, x1 = (bx + kb + 1) * sx - 1, y0 = 4 * (by - kb), y1 = 4 * (by + kb) + 3, z0 = 4 * (bz - kb), z1 = 4 * (bz + kb) + 3;
  if (x0 > 0) rc = fminf(rc, q->px - (float)x0 * cx);
  if (x1 < dx - 1) rc = fminf(rc, (float)(x1 + 1) * cx - q->px);
  if (y0 > 0) rc = fminf(rc, q->py - (float)y0 * c);
  if (y1 < dy - 1) rc = fminf(rc, (float)(y1 + 1) * c - q->py);
  if (z0 > 0) rc = fminf(rc, q->pz - (float)z0 * c);
  if (z1 < dz - 1) rc = fminf(rc, (float)(z1 + 1) * c - q->pz);
  if (rc == INFINITY) return INFINITY;
  rc = fmaxf(rc - margin, 0.0f); return rc * rc * SHRINK;
}
static void far_run(Qs* q) {
  f_n++; if (q->bd2 == INFINITY) f_nobest++;
  float b2 = bound2(q);
  int bx = (q->ix / xs) >> 2, by = q->iy >> 2, bz = q->iz >> 2;
  far_block(q, &b2, bx, by, bz, 1);
  int kmax = nbx > nby ? nbx : nby; if (nbz > kmax) kmax = nbz;
  for (int kb = 1; kb <= kmax; kb++) {
    float rc2 = far_cert(q, bx, by, bz, kb - 1);
    if (rc2 == INFINITY || b2 <= rc2) return;
    f_shells++;
    for (int Z = bz - kb; Z <= bz + kb; Z++) for (int Y = by - kb; Y <= by + kb; Y++) for (int X = bx - kb; X <= bx + kb; X++) {
      if (Z < 0 || Y < 0 || X < 0 || Z >= nbz || Y >= nby || X >= nbx) continue;
      if (abs(Z - bz) != kb && abs(Y - by) != kb && abs(X - bx) != kb) continue;
      far_block(q, &b2, X, Y, Z, 1);
    }
  }
}

typedef struct { uint32_t key, i; } KV;
static int cmpkv(const void* a, const void* b) { uint32_t x = ((const KV*)a)->key, y = ((const KV*)b)->key; return x < y ? -1 : (x > y ? 1 : (((const KV*)a)->i < ((const KV*)b)->i ? -1 : 1)); }

int main(int argc, char** argv) {
  c = argc > 1 ? atof(argv[1]) : 0.25f; xs = argc > 2 ? atoi(argv[2]) : 4; int K1 = argc > 3 ? atoi(argv[3]) : 1;
  int variant = argc > 4 ? atoi(argv[4]) : 0; int SW = argc > 5 ? atoi(argv[5]) : 0; int skipbits = argc > 6 ? atoi(argv[6]) : 4;
  size_t s; float* P = rd("pts.f32", &s); int64_t N = s / 12;
  float* Qf = rd("q.f32", &s); int64_t Q = s / 12; float* rad = rd("rad.f32", &s);
  float lo[3] = {1e30, 1e30, 1e30}, hi[3] = {-1e30, -1e30, -1e30};
  for (int64_t i = 0; i < N; i++) for (int d = 0; d < 3; d++) { lo[d] = fminf(lo[d], P[3 * i + d]); hi[d] = fmaxf(hi[d], P[3 * i + d]); }
  ox = lo[0]; oy = lo[1]; oz = lo[2]; inv_c = 1.0f / c; cx = c / xs; inv_cx = inv_c * xs;
  dx = (int)floor((hi[0] - lo[0]) * inv_cx) + 1; dy = (int)floor((hi[1] - lo[1]) * inv_c) + 1; dz = (int)floor((hi[2] - lo[2]) * inv_c) + 1;
  ncells = (int64_t)dx * dy * dz; margin = 1e-3f * cx + 8.0f * 200.0f * 1.2e-7f;
  cell_start = calloc(ncells + 2, 4); pts = malloc(sizeof(P4) * N);
  uint32_t* key = malloc(4 * N);
  for (int64_t i = 0; i < N; i++) {
    int ix = clampi((int)floorf((P[3 * i] - ox) * inv_cx), 0, dx - 1), iy = clampi((int)floorf((P[3 * i + 1] - oy) * inv_c), 0, dy - 1), iz = clampi((int)floorf((P[3 * i + 2] - oz) * inv_c), 0, dz - 1);
    key[i] = (uint32_t)(((int64_t)iz * dy + iy) * dx + ix); cell_start[key[i] + 1]++;
  }
  int64_t occ = 0; for (int64_t k = 0; k < ncells; k++) { if (cell_start[k + 1]) occ++; cell_start[k + 1] += cell_start[k]; }
  uint32_t* cur = malloc(4 * (ncells + 1)); memcpy(cur, cell_start, 4 * (ncells + 1));
  for (int64_t i = 0; i < N; i++) { uint32_t j = cur[key[i]]++; pts[j].x = P[3 * i]; pts[j].y = P[3 * i + 1]; pts[j].z = P[3 * i + 2]; pts[j].id = (uint32_t)i; }
  free(cur);
  nbx = ((dx + xs - 1) / xs + 3) / 4; nby = (dy + 3) / 4; nbz = (dz + 3) / 4;
  occw = calloc((size_t)nbx * nby * nbz, 8);
  for (int64_t i = 0; i < N; i++) { uint32_t k = key[i]; int ix = k % dx, iy = (k / dx) % dy, iz = k / ((uint32_t)dx * dy); int cxx = ix / xs;
    occw[((size_t)(iz >> 2) * nby + (iy >> 2)) * nbx + (cxx >> 2)] |= 1ull << ((((iz & 3) << 2) | (iy & 3)) << 2 | (cxx & 3)); }
  printf("grid c=%.4f xs=%d dims %d %d %d cells %lld occupied %lld (%.2f pts/occ cell) K1=%d variant=%d SW=%d\n", c, xs, dx, dy, dz, (long long)ncells, (long long)occ, (double)N / occ, K1, variant, SW);
  KV* kv = malloc(sizeof(KV) * Q);
  for (int64_t i = 0; i < Q; i++) {
    float px = Qf[3 * i] - ox, py = Qf[3 * i + 1] - oy, pz = Qf[3 * i + 2] - oz;
    int ix = clampi((int)floorf(px * inv_cx), 0, dx - 1), iy = clampi((int)floorf(py * inv_c), 0, dy - 1), iz = clampi((int)floorf(pz * inv_c), 0, dz - 1);
    if (skipbits >= 0) kv[i].key = (uint32_t)(((int64_t)iz * dy + iy) * dx + ix) >> skipbits;
    else { int T = -skipbits; int tx = ix / (T * xs), ty = iy / T, tz = iz / T; kv[i].key = (uint32_t)(((int64_t)tz * (dy / T + 1) + ty) * (dx / (T * xs) + 1) + tx); }
    kv[i].i = (uint32_t)i;
  }
  qsort(kv, Q, sizeof(KV), cmpkv);
  double tot_c = 0, tot_vis = 0, cert = 0, cost_rowmax = 0, cost_sm = 0, cost_rows_any = 0; int64_t hist[16] = {0};
  double tot_c_cert = 0;
  static Qs W[32];
  for (int64_t w0 = 0; w0 < Q; w0 += 32) {
    int nl = (int)(Q - w0 < 32 ? Q - w0 : 32), maxsite = 0;
    for (int l = 0; l < nl; l++) {
      Qs* q = &W[l]; int64_t i = kv[w0 + l].i;
      q->qx = Qf[3 * i]; q->qy = Qf[3 * i + 1]; q->qz = Qf[3 * i + 2]; q->px = q->qx - ox; q->py = q->qy - oy; q->pz = q->qz - oz;
      q->ix = clampi((int)floorf(q->px * inv_cx), 0, dx - 1); q->iy = clampi((int)floorf(q->py * inv_c), 0, dy - 1); q->iz = clampi((int)floorf(q->pz * inv_c), 0, dz - 1);
      float r = rad[i]; q->R2 = r * r * 1.00001f + 1e-12f;
      int ok = search(q, K1, variant, SW);
      if (!ok) far_run(q);
      cert += ok; int t = 0; for (int k = 0; k < q->nsite; k++) t += q->len[k];
      tot_c += t; tot_vis += q->visited; if (ok) tot_c_cert += t; if (q->nsite > maxsite) maxsite = q->nsite;
      int hb = t == 0 ? 0 : (int)log2(t) + 1; if (hb > 15) hb = 15; hist[hb]++;
    }
    { // grouped model: group A = site 0, group B = site 1 (+2 for variant 0), group C = the rest, each scanned by one flat loop
      int nA = 1, nB = (variant == 0) ? 3 : 2;
      int mA = 0, mB = 0, mC = 0;
      for (int l = 0; l < nl; l++) { int a = 0, b = 0, cc = 0; for (int k = 0; k < W[l].nsite; k++) { if (k < nA) a += W[l].len[k]; else if (k < nB) b += W[l].len[k]; else cc += W[l].len[k]; }
        if (a > mA) mA = a; if (b > mB) mB = b; if (cc > mC) mC = cc; }
      cost_grp += mA + mB + mC; cost_grpA += mA; cost_grpB += mB; cost_grpC += mC;
    }
    int msum = 0;
    for (int l = 0; l < nl; l++) { int t = 0; for (int k = 0; k < W[l].nsite; k++) t += W[l].len[k]; if (t > msum) msum = t; }
    cost_sm += msum;
    for (int k = 0; k < maxsite; k++) {
      int m = 0; for (int l = 0; l < nl; l++) if (k < W[l].nsite && W[l].len[k] > m) m = W[l].len[k];
      cost_rowmax += m; if (m) cost_rows_any++;
    }
  }
  double warps = ceil(Q / 32.0);
  printf("candidates/query %.1f  visited rows/query %.2f  certified %.4f\n", tot_c / Q, tot_vis / Q, cert / Q);
  printf("warp model: rowmax iterations/warp %.1f (lane eff %.3f), live sites/warp %.1f; state-machine iterations/warp %.1f (eff %.3f); flat %.1f\n",
         cost_rowmax / warps, tot_c / (32 * cost_rowmax), cost_rows_any / warps, cost_sm / warps, tot_c / (32 * cost_sm), tot_c / 32 / warps);
  printf("grouped model: iterations/warp %.1f (own %.1f centre %.1f ring %.1f) eff %.3f\n", cost_grp / warps, cost_grpA / warps, cost_grpB / warps, cost_grpC / warps, tot_c / (32 * cost_grp));
  printf("far: n %.0f (nobest %.0f) per query: shells %.2f blocks %.1f words %.1f nonzero %.1f bits(box) %.1f cells %.2f cands %.1f\n", f_n, f_nobest, f_shells / f_n, f_blocks / f_n, f_words / f_n, f_nzwords / f_n, f_bits / f_n, f_cells / f_n, f_cands / f_n);
  printf("hist(log2 cand):"); for (int k = 0; k < 16; k++) printf(" %lld", (long long)hist[k]); printf("\n");
  return 0;
}
