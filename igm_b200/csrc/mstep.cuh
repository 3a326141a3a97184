// mstep.cuh -- M-step device code: one CTA per structure, the structure's atoms resident in
// shared memory for the whole annealing protocol.
//
// Force field (restated from the script igm/model/kernel/lammps.py:149-358 generates and the
// coefficients lammps.py:63-146 writes; see oracle/ff_oracle.c for the pin status of each term):
//   * pair_style soft between NORMAL beads, A = ((ri+rj)/pi)^2 * evfactor * stage_scale, rc = ri+rj
//   * harmonic_upper_bound / harmonic_lower_bound bonds (polymer, Hi-C, FISH, SPRITE ...)
//   * fix ellipsoidalenvelope a b c k per envelope group (k<0: DamID lamina branch)
//   * fix setforce 0 on DUMMY_STATIC atoms
// Evaluation is gather-style: every atom sums the forces of its own neighbour / bond rows, a
// team of TEAM lanes per atom with a shuffle reduction.  No atomics => bit-reproducible runs.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>
#include "../../include/igmk.h"

namespace igmk {

constexpr int MD_THREADS   = 1024;
constexpr int TEAM         = 8;        // lanes cooperating on one atom's rows
constexpr int NL_WIDTH0    = 96;       // initial neighbour slots per bead (full list, cutoff + skin); doubled on overflow
constexpr int MAX_CELLS    = 4096;     // cell grid capacity (16^3)
constexpr int MAX_CELL_DIM = 16;

// One structure of the batch, device-side view.
struct StructDesc {
    int32_t atom_off;        // offset into per-atom arrays
    int32_t n_atoms;
    int32_t n_nonfixed;      // atoms integrated (beads + centroids)
    int32_t bond_ptr_off;    // offset into bond_ptr (n_atoms+1 entries)
    int64_t bond_item_off;   // offset into bond_items
    int32_t btype_off;       // offset into bond type tables
    int32_t n_env;
    double  env[IGMK_MAX_ENVELOPES][4];   // a, b, c, k
    double  evfactor;        // < 0: no sterics
    int64_t seed;
    float   maxrad;          // largest bead radius (skin = maxrad, lammps.py:216)
    int32_t pad;
};

struct Batch {
    int32_t n_struct;
    int32_t max_atoms;
    const StructDesc* desc;
    float4*   xyzr;          // [atoms] x, y, z, radius  (in: start coordinates, out: result)
    const uint8_t*  kind;    // [atoms]
    const uint32_t* envmask; // [atoms] bit e set: atom belongs to envelope e
    const int32_t*  bond_ptr;    // CSR rows by atom (directional: each bond appears in both rows)
    const uint32_t* bond_items;  // partner | type << 16
    const float2*   btype_f;     // (r0, k) ; k < 0 encodes a lower bound
    const double2*  btype_d;
    uint16_t* nl;            // [atoms][nl_width]
    int32_t   nl_width;
    uint16_t* nl_cnt;        // [atoms]
    double*   xd;            // [atoms*3] double coordinates (CG)
    double*   fd;            // [atoms*3] double forces (single point / CG output)
    double*   work;          // [atoms*12] CG vectors h, g, x0 + neighbour-build coordinates
    double*   big;           // [(atoms + 32*n_struct)*10] state of structures too large for shared memory (or NULL)
    igmk_result_t* res;      // [n_struct]
    int32_t*  flags;         // [n_struct] bit0: neighbour overflow, bit1: non-finite
};

// ---------------------------------------------------------------------------------------
template <typename T> struct Num;
template <> struct Num<float> {
    using vec4 = float4;
    static __device__ __forceinline__ float sqrt_(float x) { return sqrtf(x); }
    static __device__ __forceinline__ void sincospi_(float x, float* s, float* c) { sincospif(x, s, c); }
    static __device__ __forceinline__ float sinpi_(float x) { return sinpif(x); }
    static __device__ __forceinline__ float abs_(float x) { return fabsf(x); }
};
template <> struct Num<double> {
    using vec4 = double4;
    static __device__ __forceinline__ double sqrt_(double x) { return sqrt(x); }
    static __device__ __forceinline__ void sincospi_(double x, double* s, double* c) { sincospi(x, s, c); }
    static __device__ __forceinline__ double sinpi_(double x) { return sinpi(x); }
    static __device__ __forceinline__ double abs_(double x) { return fabs(x); }
};

template <typename T>
__device__ __forceinline__ T warp_sum(T v)
{
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
template <typename T>
__device__ __forceinline__ T warp_max(T v)
{
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) { T w = __shfl_xor_sync(0xffffffffu, v, o); v = w > v ? w : v; }
    return v;
}
template <typename T>
__device__ __forceinline__ T warp_min(T v)
{
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) { T w = __shfl_xor_sync(0xffffffffu, v, o); v = w < v ? w : v; }
    return v;
}

// Deterministic block-wide sum of NV values: shuffle tree per warp, then every thread adds the
// warp partials in warp order.  `buf` must hold 32*NV entries and be free for reuse by the caller
// only after the next barrier (callers alternate two buffers).
template <typename T, int NV>
__device__ __forceinline__ void block_sum(T (&v)[NV], T* buf)
{
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
#pragma unroll
    for (int k = 0; k < NV; k++) v[k] = warp_sum(v[k]);
    if (lane == 0) {
#pragma unroll
        for (int k = 0; k < NV; k++) buf[warp * NV + k] = v[k];
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < NV; k++) {
        T s = 0;
        for (int w = 0; w < nw; w++) s += buf[w * NV + k];
        v[k] = s;
    }
}

// Park-Miller generator of LAMMPS' RanPark (`velocity create`): state after `n` steps from `seed`.
__device__ __forceinline__ uint32_t ranpark_jump(uint32_t seed, uint32_t n)
{
    const uint64_t M = 2147483647ull;
    uint64_t base = 16807ull, acc = seed % M;
    while (n) {
        if (n & 1u) acc = (acc * base) % M;
        base = (base * base) % M;
        n >>= 1;
    }
    return (uint32_t)acc;
}

// ---------------------------------------------------------------------------------------
// Per-CTA view of one structure.  T = float for the MD loop, double for single points and CG.
template <typename T>
struct Ctx {
    using V4 = typename Num<T>::vec4;
    int n;                       // atoms
    int n_nonfixed;
    V4* x;                       // smem: x, y, z, radius
    T *fx, *fy, *fz;             // smem forces
    T *bx, *by, *bz;             // smem: coordinates at the last neighbour build
    uint16_t* cell_start;        // smem [MAX_CELLS+1]
    uint16_t* cell_atoms;        // smem [n] atoms sorted by cell
    uint16_t* cell_of;           // smem [n]
    int*      ibuf;              // smem small int scratch [64]
    T*        rbuf;              // smem reduction scratch [2][32*8]
    T*        envp;              // smem [IGMK_MAX_ENVELOPES][4]: semi-axes * stage scale, k
    const uint8_t*  kind;
    const uint32_t* envmask;
    const int32_t*  bptr;
    const uint32_t* bitems;
    const void*     btype;       // float2* or double2*
    uint16_t* nl;
    uint16_t* nl_cnt;
    int nlw;                     // neighbour slots per bead
    const StructDesc* d;
    T ev_pref;                   // evfactor * stage scale / pi^2
    T bond_pref;
    T skin;
    int n_rebuild;
    int overflow;
    int rb_parity;

    __device__ __forceinline__ T* red() { rb_parity ^= 1; return rbuf + rb_parity * 256; }

    // per-stage force-field scales: pair prefactor (fix adapt, lammps.py:287-290) and envelope
    // semi-axes (lammps.py:298-299).  Ends with a barrier.
    __device__ void set_stage(double ev_scale, double env_scale)
    {
        const double PI2 = 3.14159265358979323846 * 3.14159265358979323846;
        ev_pref = d->evfactor < 0 ? T(-1) : T(d->evfactor * ev_scale / PI2);
        if (threadIdx.x < IGMK_MAX_ENVELOPES * 4) {
            const int e = threadIdx.x >> 2, k = threadIdx.x & 3;
            envp[threadIdx.x] = e < d->n_env ? T(k < 3 ? d->env[e][k] * env_scale : d->env[e][3]) : T(0);
        }
        __syncthreads();
    }

    // ---- neighbour list: cell binning in shared memory, full list with skin -------------
    __device__ void rebuild()
    {
        const int tid = threadIdx.x, nt = blockDim.x;
        // bounding box of the beads
        T lo[3] = {T(1e30), T(1e30), T(1e30)}, hi[3] = {T(-1e30), T(-1e30), T(-1e30)};
        for (int i = tid; i < n; i += nt) {
            V4 p = x[i];
            bx[i] = p.x; by[i] = p.y; bz[i] = p.z;
            if (kind[i] == IGMK_KIND_BEAD) {
                lo[0] = p.x < lo[0] ? p.x : lo[0]; hi[0] = p.x > hi[0] ? p.x : hi[0];
                lo[1] = p.y < lo[1] ? p.y : lo[1]; hi[1] = p.y > hi[1] ? p.y : hi[1];
                lo[2] = p.z < lo[2] ? p.z : lo[2]; hi[2] = p.z > hi[2] ? p.z : hi[2];
            }
        }
        T* rb = red();
        {
            const int lane = tid & 31, warp = tid >> 5, nw = (nt + 31) >> 5;
#pragma unroll
            for (int k = 0; k < 3; k++) { lo[k] = warp_min(lo[k]); hi[k] = warp_max(hi[k]); }
            if (lane == 0) {
#pragma unroll
                for (int k = 0; k < 3; k++) { rb[warp * 6 + k] = lo[k]; rb[warp * 6 + 3 + k] = hi[k]; }
            }
            for (int c = tid; c <= MAX_CELLS; c += nt) cell_start[c] = 0;
            __syncthreads();
            for (int w = 0; w < nw; w++) {
#pragma unroll
                for (int k = 0; k < 3; k++) {
                    T a = rb[w * 6 + k], b = rb[w * 6 + 3 + k];
                    lo[k] = a < lo[k] ? a : lo[k]; hi[k] = b > hi[k] ? b : hi[k];
                }
            }
        }
        const T rlist = T(2) * T(d->maxrad) + skin;
        int dim[3]; T inv[3];
#pragma unroll
        for (int k = 0; k < 3; k++) {
            T ext = hi[k] - lo[k];
            if (!(ext > T(0))) ext = T(0);
            int m = (int)(ext / rlist) + 1;
            if (m > MAX_CELL_DIM) m = MAX_CELL_DIM;
            T cw = ext / T(m);
            if (cw < rlist) cw = rlist;
            dim[k] = m; inv[k] = T(1) / cw;
        }
        const int ncell = dim[0] * dim[1] * dim[2];
        // count atoms per cell (integer shared-memory atomics: counts are order independent)
        for (int i = tid; i < n; i += nt) {
            int c = 0xffff;
            if (kind[i] == IGMK_KIND_BEAD) {
                V4 p = x[i];
                int cx = (int)((p.x - lo[0]) * inv[0]), cy = (int)((p.y - lo[1]) * inv[1]), cz = (int)((p.z - lo[2]) * inv[2]);
                cx = min(max(cx, 0), dim[0] - 1); cy = min(max(cy, 0), dim[1] - 1); cz = min(max(cz, 0), dim[2] - 1);
                c = (cz * dim[1] + cy) * dim[0] + cx;
                // 16-bit counters packed two per word
                atomicAdd((unsigned int*)cell_start + ((c + 1) >> 1), ((c + 1) & 1) ? 0x10000u : 1u);
            }
            cell_of[i] = (uint16_t)c;
        }
        __syncthreads();
        // exclusive scan of cell_start[1..ncell] in place (cell_start[c+1] held the count of cell c)
        {
            const int per = (ncell + nt - 1) / nt;
            const int c0 = tid * per, c1 = min(c0 + per, ncell);
            int s = 0;
            for (int c = c0; c < c1; c++) s += cell_start[c + 1];
            // block exclusive scan of s
            const int lane = tid & 31, warp = tid >> 5, nw = (nt + 31) >> 5;
            int incl = s;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) { int t = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += t; }
            if (lane == 31) ibuf[warp] = incl;
            __syncthreads();
            int base = 0;
            for (int w = 0; w < warp; w++) base += ibuf[w];
            (void)nw;
            int run = base + incl - s;
            for (int c = c0; c < c1; c++) { int cnt = cell_start[c + 1]; cell_start[c + 1] = (uint16_t)(run + cnt); run += cnt; }
            // now cell_start[c+1] = inclusive end of cell c; cell_start[0] = 0
        }
        __syncthreads();
        // stable fill: one thread per cell collects its atoms in index order would be O(n) per
        // cell; instead scatter with a per-cell cursor and sort the (tiny) cells afterwards.
        // cursors: reuse nl_cnt-independent scratch -> use bx as int storage is unsafe; use cell_atoms
        // two-step: every atom finds its rank inside its cell by counting lower-indexed atoms of the
        // same cell among the cell's final members, so first drop atoms unordered via atomics ...
        // cursors live behind the atom list, from an EVEN offset: with an odd n the 32-bit word of cursor 0 would also hold
        // atom slot n-1, which other threads write with plain 16-bit stores in this very phase
        const int cur0 = (n + 1) & ~1;
        for (int c = tid; c < ncell; c += nt) cell_atoms[cur0 + c] = 0;
        __syncthreads();
        for (int i = tid; i < n; i += nt) {
            int c = cell_of[i];
            if (c == 0xffff) continue;
            // 16-bit cursor increments through a 32-bit CAS-free atomicAdd on the containing word
            unsigned int* w = (unsigned int*)(cell_atoms + ((cur0 + c) & ~1));
            unsigned int old = atomicAdd(w, ((cur0 + c) & 1) ? 0x10000u : 1u);
            int pos = ((cur0 + c) & 1) ? (old >> 16) : (old & 0xffffu);
            cell_atoms[cell_start[c] + pos] = (uint16_t)i;
        }
        __syncthreads();
        // ... then insertion-sort each cell by atom index so that the list order (and with it the
        // floating-point summation order) does not depend on the race above
        for (int c = tid; c < ncell; c += nt) {
            int b = cell_start[c], e = cell_start[c + 1];
            for (int p = b + 1; p < e; p++) {
                uint16_t v = cell_atoms[p]; int q = p - 1;
                while (q >= b && cell_atoms[q] > v) { cell_atoms[q + 1] = cell_atoms[q]; q--; }
                cell_atoms[q + 1] = v;
            }
        }
        __syncthreads();
        // neighbour search, a team per bead
        const int team = tid / TEAM, tl = tid % TEAM, nteam = nt / TEAM;
        const unsigned tmask = ((1u << TEAM) - 1u) << ((tid & 31) / TEAM * TEAM);
        int ovf = 0;
        for (int base = 0; base < n; base += nteam) {
            const int i = base + team;
            const bool valid = i < n && cell_of[min(i, n - 1)] != 0xffff;
            int cnt = 0;
            if (valid) {
                const V4 p = x[i];
                const int c = cell_of[i];
                const int cx = c % dim[0], cy = (c / dim[0]) % dim[1], cz = c / (dim[0] * dim[1]);
                uint16_t* row = nl + (size_t)i * nlw;
                for (int dz = -1; dz <= 1; dz++) {
                    const int az = cz + dz; if (az < 0 || az >= dim[2]) continue;
                    for (int dy = -1; dy <= 1; dy++) {
                        const int ay = cy + dy; if (ay < 0 || ay >= dim[1]) continue;
                        // the three x-neighbour cells are contiguous in the sorted atom array
                        const int ax0 = max(cx - 1, 0), ax1 = min(cx + 1, dim[0] - 1);
                        const int cb = (az * dim[1] + ay) * dim[0];
                        const int b = cell_start[cb + ax0], e = cell_start[cb + ax1 + 1];
                        for (int q0 = b; q0 < e; q0 += TEAM) {
                            const int q = q0 + tl;
                            bool hit = false; int j = 0;
                            if (q < e) {
                                j = cell_atoms[q];
                                const V4 pj = x[j];
                                const T ddx = p.x - pj.x, ddy = p.y - pj.y, ddz = p.z - pj.z;
                                const T rl = p.w + pj.w + skin;
                                hit = (j != i) && (ddx * ddx + ddy * ddy + ddz * ddz < rl * rl);
                            }
                            const unsigned bal = __ballot_sync(tmask, hit) & tmask;
                            if (hit) {
                                const int pos = cnt + __popc(bal & ((1u << (tid & 31)) - 1u));
                                if (pos < nlw) row[pos] = (uint16_t)j;
                            }
                            cnt += __popc(bal);
                        }
                    }
                }
                if (cnt > nlw) { ovf = 1; cnt = nlw; }
                if (tl == 0) nl_cnt[i] = (uint16_t)cnt;
            } else if (i < n && tl == 0) {
                nl_cnt[i] = 0;
            }
        }
        overflow |= __syncthreads_or(ovf);
        n_rebuild++;
    }

    // ---- force on atom i, evaluated by a team; valid on team lane 0 after the reduction ----
    template <bool ENERGY>
    __device__ __forceinline__ void atom_force(int i, int tl, unsigned tmask, T& ofx, T& ofy, T& ofz,
                                               T& e_pair, T& e_bond, T* e_env)
    {
        const V4 p = x[i];
        T ax = 0, ay = 0, az = 0;
        const uint8_t kd = kind[i];
        if (kd == IGMK_KIND_BEAD && ev_pref >= T(0)) {
            const int cnt = nl_cnt[i];
            const uint16_t* row = nl + (size_t)i * nlw;
            for (int k = tl; k < cnt; k += TEAM) {
                const int j = row[k];
                const V4 pj = x[j];
                const T dx = p.x - pj.x, dy = p.y - pj.y, dz = p.z - pj.z;
                const T rc = p.w + pj.w;
                const T r2 = dx * dx + dy * dy + dz * dz;
                if (r2 < rc * rc) {
                    const T r = Num<T>::sqrt_(r2);
                    const T irc = T(1) / rc;
                    const T A = rc * rc * ev_pref;
                    T s, c;
                    if (ENERGY) { Num<T>::sincospi_(r * irc, &s, &c); e_pair += T(0.5) * A * (T(1) + c); }
                    else s = Num<T>::sinpi_(r * irc);
                    if (r > T(0)) {
                        const T fp = A * s * T(3.14159265358979323846) * irc / r;
                        ax += dx * fp; ay += dy * fp; az += dz * fp;
                    }
                }
            }
        }
        // a frozen atom's own row only matters for the energy (its force is zeroed)
        if (ENERGY || kd != IGMK_KIND_STATIC) {
            const int b0 = bptr[i], b1 = bptr[i + 1];
            for (int q = b0 + tl; q < b1; q += TEAM) {
                const uint32_t it = bitems[q];
                const int j = it & 0xffffu;
                T r0, k;
                if (sizeof(T) == 4) { float2 t = ((const float2*)btype)[it >> 16]; r0 = t.x; k = t.y; }
                else { double2 t = ((const double2*)btype)[it >> 16]; r0 = t.x; k = t.y; }
                const V4 pj = x[j];
                const T dx = p.x - pj.x, dy = p.y - pj.y, dz = p.z - pj.z;
                const T r2 = dx * dx + dy * dy + dz * dz;
                const bool act = k > T(0) ? (r2 > r0 * r0) : (r2 < r0 * r0);
                if (act) {
                    const T r = Num<T>::sqrt_(r2);
                    const T dr = r - r0;
                    const T rk = bond_pref * Num<T>::abs_(k) * dr;
                    if (ENERGY) e_bond += T(0.5) * rk * dr;
                    if (r > T(0)) {
                        const T fb = T(-2) * rk / r;
                        ax += dx * fb; ay += dy * fb; az += dz * fb;
                    }
                }
            }
        }
#pragma unroll
        for (int o = TEAM / 2; o > 0; o >>= 1) {
            ax += __shfl_xor_sync(tmask, ax, o);
            ay += __shfl_xor_sync(tmask, ay, o);
            az += __shfl_xor_sync(tmask, az, o);
        }
        if (tl == 0 && kd != IGMK_KIND_STATIC) {
            uint32_t m = envmask[i];
            while (m) {
                const int e = __ffs(m) - 1; m &= m - 1;
                const T a = envp[4 * e] - p.w, b = envp[4 * e + 1] - p.w, c = envp[4 * e + 2] - p.w, k = envp[4 * e + 3];
                const T k2 = p.x * p.x / (a * a) + p.y * p.y / (b * b) + p.z * p.z / (c * c);
                const bool act = k > T(0) ? (k2 > T(1)) : (k < T(0) && k2 < T(1));
                const T n2 = p.x * p.x + p.y * p.y + p.z * p.z;
                if (act && n2 > T(0)) {
                    const T g = T(1) - T(1) / Num<T>::sqrt_(k2);
                    const T ak = Num<T>::abs_(k);
                    if (ENERGY) e_env[e] += T(0.5) * ak * g * g * n2;
                    ax -= ak * g * p.x; ay -= ak * g * p.y; az -= ak * g * p.z;
                }
            }
        }
        if (kd == IGMK_KIND_STATIC) { ax = 0; ay = 0; az = 0; }
        ofx = ax; ofy = ay; ofz = az;
    }

    // forces for all atoms into fx/fy/fz (+ optional energies, block-reduced into e[0..2+n_env])
    template <bool ENERGY>
    __device__ void forces(T* e_out)
    {
        const int tid = threadIdx.x, nt = blockDim.x;
        const int team = tid / TEAM, tl = tid % TEAM, nteam = nt / TEAM;
        const unsigned tmask = ((1u << TEAM) - 1u) << ((tid & 31) / TEAM * TEAM);
        T ep = 0, eb = 0, ee[IGMK_MAX_ENVELOPES];
        if (ENERGY) {
#pragma unroll
            for (int k = 0; k < IGMK_MAX_ENVELOPES; k++) ee[k] = 0;
        }
        for (int base = 0; base < n; base += nteam) {
            const int i = base + team;
            if (i < n) {
                T ax, ay, az;
                atom_force<ENERGY>(i, tl, tmask, ax, ay, az, ep, eb, ee);
                if (tl == 0) { fx[i] = ax; fy[i] = ay; fz[i] = az; }
            }
        }
        if (ENERGY) {
            T v[2 + IGMK_MAX_ENVELOPES];
            v[0] = ep; v[1] = eb;
#pragma unroll
            for (int k = 0; k < IGMK_MAX_ENVELOPES; k++) v[2 + k] = ee[k];
            // two reductions of <= 8 values each (scratch holds 8 per warp)
            T a[8];
#pragma unroll
            for (int k = 0; k < 8; k++) a[k] = v[k];
            block_sum<T, 8>(a, red());
            T b2[2];
            b2[0] = v[8]; b2[1] = v[9];
            __syncthreads();
            block_sum<T, 2>(b2, red());
#pragma unroll
            for (int k = 0; k < 8; k++) e_out[k] = a[k];
            e_out[8] = b2[0]; e_out[9] = b2[1];
        }
        __syncthreads();
    }

    // true if some atom moved more than skin/2 since the last build (block-uniform result)
    __device__ bool stale()
    {
        const T lim = T(0.25) * skin * skin;
        int need = 0;
        for (int i = threadIdx.x; i < n; i += blockDim.x) {
            V4 p = x[i];
            T dx = p.x - bx[i], dy = p.y - by[i], dz = p.z - bz[i];
            need |= (dx * dx + dy * dy + dz * dz > lim);
        }
        return __syncthreads_or(need) != 0;
    }
};

// shared-memory carve-up -----------------------------------------------------------------
// float: everything in shared memory.  double (single point / CG): the coordinates of the last
// neighbour build go to global memory so that 2 Mb diploid structures still fit one CTA.
// Structures too large for that (ctx_is_big: > ~3500 atoms in double, e.g. the 30 k beads of a 200 kb
// genome) keep positions, forces and build coordinates in a global scratch area of 10 values per
// atom instead; only the cell grid and the reduction scratch stay in shared memory.
constexpr size_t CTX_SMEM_LIMIT = 227 * 1024;

template <typename T>
__host__ __device__ inline size_t ctx_smem_fixed(int n_atoms)
{
    const size_t np = (size_t)((n_atoms + 31) / 32 * 32);
    size_t b = 0;
    b += 2 * 256 * sizeof(T);                // reduction scratch
    b += IGMK_MAX_ENVELOPES * 4 * sizeof(T); // envelope parameters
    b += 64 * sizeof(int);                   // int scratch
    b += ((size_t)MAX_CELLS + 2) * 2;        // cell_start
    b += (np + MAX_CELLS + 2) * 2;           // cell_atoms + cursors
    b += np * 2;                             // cell_of
    return b;
}

template <typename T>
__host__ __device__ inline size_t ctx_smem_state(int n_atoms, bool with_velocity)
{
    const size_t np = (size_t)((n_atoms + 31) / 32 * 32);
    size_t b = np * 4 * sizeof(T) + np * 3 * sizeof(T);      // x, f
    if (sizeof(T) == 4) b += np * 3 * sizeof(T);             // build coordinates
    if (with_velocity) b += np * 3 * sizeof(T);
    return b;
}

template <typename T>
__host__ __device__ inline bool ctx_is_big(int n_atoms, bool with_velocity)
{
    return (ctx_smem_fixed<T>(n_atoms) + ctx_smem_state<T>(n_atoms, with_velocity) + 255) / 256 * 256 > CTX_SMEM_LIMIT;
}

template <typename T>
__host__ __device__ inline size_t ctx_smem_bytes(int n_atoms, bool with_velocity)
{
    size_t b = ctx_smem_fixed<T>(n_atoms);
    if (!ctx_is_big<T>(n_atoms, with_velocity)) b += ctx_smem_state<T>(n_atoms, with_velocity);
    return (b + 255) / 256 * 256;
}

// gbig: this structure's global scratch (10 doubles per atom, Batch::big), used in big mode only
template <typename T>
__device__ inline T* ctx_carve(Ctx<T>& c, unsigned char* smem, int n_atoms, bool with_velocity, T* gbuild, double* gbig = nullptr)
{
    const size_t np = (size_t)((n_atoms + 31) / 32 * 32);
    unsigned char* p = smem;
    T* vel = nullptr;
    if (ctx_is_big<T>(n_atoms, with_velocity)) {
        T* g = (T*)gbig;
        c.x = (typename Ctx<T>::V4*)g; g += np * 4;
        c.fx = g; c.fy = c.fx + np; c.fz = c.fy + np; g += np * 3;
        c.bx = g; c.by = c.bx + np; c.bz = c.by + np;
    } else {
        c.x = (typename Ctx<T>::V4*)p; p += np * 4 * sizeof(T);
        c.fx = (T*)p; c.fy = c.fx + np; c.fz = c.fy + np; p += np * 3 * sizeof(T);
        if (sizeof(T) == 4) { c.bx = (T*)p; c.by = c.bx + np; c.bz = c.by + np; p += np * 3 * sizeof(T); }
        else { c.bx = gbuild; c.by = gbuild + n_atoms; c.bz = gbuild + 2 * (size_t)n_atoms; }
        if (with_velocity) { vel = (T*)p; p += np * 3 * sizeof(T); }
    }
    c.rbuf = (T*)p; p += 2 * 256 * sizeof(T);
    c.envp = (T*)p; p += IGMK_MAX_ENVELOPES * 4 * sizeof(T);
    c.ibuf = (int*)p; p += 64 * sizeof(int);
    c.cell_start = (uint16_t*)p; p += ((size_t)MAX_CELLS + 2) * 2;
    c.cell_atoms = (uint16_t*)p; p += (np + MAX_CELLS + 2) * 2;
    c.cell_of = (uint16_t*)p;
    return vel;
}

template <typename T>
__device__ inline void ctx_bind(Ctx<T>& c, const Batch& B, const StructDesc* d)
{
    c.d = d;
    c.n = d->n_atoms; c.n_nonfixed = d->n_nonfixed;
    c.kind = B.kind + d->atom_off;
    c.envmask = B.envmask + d->atom_off;
    c.bptr = B.bond_ptr + d->bond_ptr_off;
    c.bitems = B.bond_items + d->bond_item_off;
    c.btype = sizeof(T) == 4 ? (const void*)(B.btype_f + d->btype_off) : (const void*)(B.btype_d + d->btype_off);
    c.nlw = B.nl_width;
    c.nl = B.nl + (size_t)d->atom_off * B.nl_width;
    c.nl_cnt = B.nl_cnt + d->atom_off;
    c.skin = T(d->maxrad);
    c.n_rebuild = 0; c.overflow = 0; c.rb_parity = 0;
}

// ---------------------------------------------------------------------------------------
// `run 0`: energies + forces at the submitted coordinates.
template <typename T>
__global__ void __launch_bounds__(MD_THREADS, 1)
k_single_point(Batch B, double ev_scale, double env_scale, double bond_pref)
{
    extern __shared__ __align__(16) unsigned char smem[];
    const StructDesc* d = B.desc + blockIdx.x;
    Ctx<T> c;
    ctx_carve<T>(c, smem, d->n_atoms, false, (T*)(B.work + (size_t)d->atom_off * 12 + (size_t)d->n_atoms * 9),
                 B.big ? B.big + ((size_t)d->atom_off + 32 * (size_t)blockIdx.x) * 10 : nullptr);
    ctx_bind(c, B, d);
    c.set_stage(ev_scale, env_scale);
    c.bond_pref = T(bond_pref);
    const float4* xin = B.xyzr + d->atom_off;
    for (int i = threadIdx.x; i < c.n; i += blockDim.x) {
        float4 p = xin[i];
        typename Ctx<T>::V4 q; q.x = p.x; q.y = p.y; q.z = p.z; q.w = p.w;
        c.x[i] = q;
    }
    __syncthreads();
    c.rebuild();
    T e[10];
    c.template forces<true>(e);
    double* f = B.fd + (size_t)d->atom_off * 3;
    for (int i = threadIdx.x; i < c.n; i += blockDim.x) {
        f[3 * i] = c.fx[i]; f[3 * i + 1] = c.fy[i]; f[3 * i + 2] = c.fz[i];
    }
    if (threadIdx.x == 0) {
        igmk_result_t* r = B.res + blockIdx.x;
        double tot = (double)e[0] + (double)e[1];
        r->pair_energy = e[0]; r->bond_energy = e[1];
        r->n_env = d->n_env;
        for (int k = 0; k < IGMK_MAX_ENVELOPES; k++) { r->env_energy[k] = k < d->n_env ? (double)e[2 + k] : 0.0; if (k < d->n_env) tot += e[2 + k]; }
        r->final_energy = tot;
        r->n_rebuilds = c.n_rebuild;
        r->status = c.overflow ? 2 : 0;
        B.flags[blockIdx.x] = c.overflow;
    }
}

// ---------------------------------------------------------------------------------------
// The annealing protocol (lammps.py:285-351): for every stage [relax run] + main run, each run =
// `velocity create` + nve/limit + temp/rescale.  float32 arithmetic, everything in shared memory.
struct MdParams {
    igmk_protocol_t prot;
    float skin_factor;       // neighbour skin = skin_factor * largest bead radius.  The LAMMPS script uses 1
                             // (`neighbor maxrad bin`, lammps.py:216); results do not depend on it, only the cost
    int   debug_stage;       // k_md2: >= 0: evaluate the forces of that stage's run setup once, store them and return
    int   general_pairs;     // k_md2: 1 = never take the uniform-radius pair loop (IGMK_GENERAL_PAIRS; tests of the general one)
};

// ---------------------------------------------------------------------------------------
// min_style cg / minimize etol ftol maxiter maxeval (lammps.py:354-355): Polak-Ribiere with the
// upstream quadratic line search, double precision.  h, g, x0 live in global memory (L2).
__global__ void __launch_bounds__(MD_THREADS, 1)
k_cg(Batch B, igmk_protocol_t pr)
{
    typedef double T;
    extern __shared__ __align__(16) unsigned char smem[];
    const StructDesc* d = B.desc + blockIdx.x;
    Ctx<T> c;
    ctx_carve<T>(c, smem, d->n_atoms, false, B.work + (size_t)d->atom_off * 12 + (size_t)d->n_atoms * 9,
                 B.big ? B.big + ((size_t)d->atom_off + 32 * (size_t)blockIdx.x) * 10 : nullptr);
    ctx_bind(c, B, d);
    const int n = c.n, tid = threadIdx.x, nt = blockDim.x;
    // unfix exVolAdapt restored A (scale 1); the last stage's envelope fix stays defined
    c.set_stage(1.0, pr.n_stage > 0 ? pr.envfactor[pr.n_stage - 1] : 1.0);
    c.bond_pref = pr.bond_prefactor;
    float4* xg = B.xyzr + d->atom_off;
    double* h = B.work + (size_t)d->atom_off * 12;
    double* g = h + (size_t)n * 3;
    double* x0 = g + (size_t)n * 3;
    for (int i = tid; i < n; i += nt) {
        float4 p = xg[i];
        double4 q; q.x = p.x; q.y = p.y; q.z = p.z; q.w = p.w;
        c.x[i] = q;
    }
    __syncthreads();
    c.rebuild();

    const double ALPHA_MAX = 1.0, ALPHA_REDUCE = 0.5, BACKTRACK_SLOPE = 0.4, QUADRATIC_TOL = 0.1,
                 EMACH = 1.0e-8, EPS_QUAD = 1.0e-28, EPS_ENERGY = 1.0e-8, DMAX = 0.1;
    double e[10];
    auto energy = [&]() -> double {
        if (c.stale()) c.rebuild();
        c.template forces<true>(e);
        double t = e[0] + e[1];
        for (int k = 0; k < d->n_env; k++) t += e[2 + k];
        return t;
    };
    auto setx = [&](double alpha) {
        for (int i = tid; i < n; i += nt) {
            double4 p = c.x[i];
            p.x = x0[3 * i] + alpha * h[3 * i]; p.y = x0[3 * i + 1] + alpha * h[3 * i + 1]; p.z = x0[3 * i + 2] + alpha * h[3 * i + 2];
            c.x[i] = p;
        }
        __syncthreads();
    };
    double ecur = energy();
    int neval = 0, iter = 0, stop = IGMK_CG_MAXITER;
    double gg;
    {
        double v[1] = {0};
        for (int i = tid; i < n; i += nt) {
            h[3 * i] = g[3 * i] = c.fx[i]; h[3 * i + 1] = g[3 * i + 1] = c.fy[i]; h[3 * i + 2] = g[3 * i + 2] = c.fz[i];
            v[0] += c.fx[i] * c.fx[i] + c.fy[i] * c.fy[i] + c.fz[i] * c.fz[i];
        }
        block_sum<T, 1>(v, c.red());
        gg = v[0];
    }
    const int nv = 3 * n;
    if (!pr.skip_minimize)
    for (iter = 0; iter < pr.max_cg_iter; iter++) {
        const double eprev = ecur;
        int fail = -1;
        {
            const double eorig = ecur;
            double v[1] = {0}; double hm = 0;
            for (int i = tid; i < n; i += nt) {
                v[0] += c.fx[i] * h[3 * i] + c.fy[i] * h[3 * i + 1] + c.fz[i] * h[3 * i + 2];
                hm = fmax(hm, fmax(fabs(h[3 * i]), fmax(fabs(h[3 * i + 1]), fabs(h[3 * i + 2]))));
            }
            block_sum<T, 1>(v, c.red());
            const double fdoth = v[0];
            // block max of hm
            hm = warp_max(hm);
            double* rb = c.red();
            if ((tid & 31) == 0) rb[tid >> 5] = hm;
            __syncthreads();
            double hmax = 0;
            for (int w = 0; w < (nt + 31) / 32; w++) hmax = fmax(hmax, rb[w]);
            if (fdoth <= 0.0) fail = IGMK_CG_DOWNHILL;
            else if (hmax == 0.0) fail = IGMK_CG_ZEROFORCE;
            else {
                const double alphamax = fmin(ALPHA_MAX, DMAX / hmax);
                for (int i = tid; i < n; i += nt) {
                    double4 p = c.x[i];
                    x0[3 * i] = p.x; x0[3 * i + 1] = p.y; x0[3 * i + 2] = p.z;
                }
                __syncthreads();
                double alpha = alphamax, fhprev = fdoth, alphaprev = 0.0, engprev = eorig;
                for (;;) {
                    setx(alpha);
                    ecur = energy(); neval++;
                    double w[1] = {0};
                    for (int i = tid; i < n; i += nt)
                        w[0] += c.fx[i] * h[3 * i] + c.fy[i] * h[3 * i + 1] + c.fz[i] * h[3 * i + 2];
                    block_sum<T, 1>(w, c.red());
                    const double fh = w[0];
                    const double delfh = fh - fhprev;
                    if (fabs(fh) < EPS_QUAD || fabs(delfh) < EPS_QUAD) {
                        setx(0.0); ecur = energy(); neval++;
                        fail = IGMK_CG_ZEROQUAD; break;
                    }
                    const double relerr = fabs(1.0 - (0.5 * (alpha - alphaprev) * (fh + fhprev) + ecur) / engprev);
                    const double alpha0 = alpha - (alpha - alphaprev) * fh / delfh;
                    if (relerr <= QUADRATIC_TOL && alpha0 > 0.0 && alpha0 < alphamax) {
                        setx(alpha0);
                        ecur = energy(); neval++;
                        if (ecur - eorig < EMACH) { fail = -1; break; }
                    }
                    const double de_ideal = -BACKTRACK_SLOPE * alpha * fdoth;
                    const double de = ecur - eorig;
                    if (de <= de_ideal) { fail = -1; break; }
                    fhprev = fh; engprev = ecur; alphaprev = alpha;
                    alpha *= ALPHA_REDUCE;
                    if (alpha <= 0.0 || de_ideal >= -EMACH) {
                        setx(0.0); ecur = energy(); neval++;
                        fail = IGMK_CG_ZEROALPHA; break;
                    }
                }
            }
        }
        if (fail >= 0) { stop = fail; break; }
        if (neval >= pr.max_cg_eval) { stop = IGMK_CG_MAXEVAL; iter++; break; }
        if (fabs(ecur - eprev) < pr.etol * 0.5 * (fabs(ecur) + fabs(eprev) + EPS_ENERGY)) { stop = IGMK_CG_ETOL; iter++; break; }
        double dd[2] = {0, 0};
        for (int i = tid; i < n; i += nt) {
            dd[0] += c.fx[i] * c.fx[i] + c.fy[i] * c.fy[i] + c.fz[i] * c.fz[i];
            dd[1] += c.fx[i] * g[3 * i] + c.fy[i] * g[3 * i + 1] + c.fz[i] * g[3 * i + 2];
        }
        block_sum<T, 2>(dd, c.red());
        if (pr.ftol > 0.0 && dd[0] < pr.ftol * pr.ftol) { stop = IGMK_CG_FTOL; iter++; break; }
        double beta = fmax(0.0, (dd[0] - dd[1]) / gg);
        if ((iter + 2) % nv == 0) beta = 0.0;
        gg = dd[0];
        double gh[1] = {0};
        for (int i = tid; i < n; i += nt) {
#pragma unroll
            for (int k = 0; k < 3; k++) {
                const double fk = k == 0 ? c.fx[i] : (k == 1 ? c.fy[i] : c.fz[i]);
                g[3 * i + k] = fk;
                const double hk = fk + beta * h[3 * i + k];
                h[3 * i + k] = hk;
                gh[0] += fk * hk;
            }
        }
        block_sum<T, 1>(gh, c.red());
        if (gh[0] <= 0.0) {
            for (int i = tid; i < n; i += nt) { h[3 * i] = g[3 * i]; h[3 * i + 1] = g[3 * i + 1]; h[3 * i + 2] = g[3 * i + 2]; }
        }
        __syncthreads();
    }
    int bad = 0;
    double* fo = B.fd + (size_t)d->atom_off * 3;
    double* xo = B.xd + (size_t)d->atom_off * 3;
    for (int i = tid; i < n; i += nt) {
        double4 p = c.x[i];
        xg[i] = make_float4((float)p.x, (float)p.y, (float)p.z, (float)p.w);
        xo[3 * i] = p.x; xo[3 * i + 1] = p.y; xo[3 * i + 2] = p.z;
        fo[3 * i] = c.fx[i]; fo[3 * i + 1] = c.fy[i]; fo[3 * i + 2] = c.fz[i];
        bad |= !(isfinite(p.x) && isfinite(p.y) && isfinite(p.z));
    }
    bad = __syncthreads_or(bad);
    if (tid == 0) {
        igmk_result_t* r = B.res + blockIdx.x;
        r->final_energy = ecur; r->pair_energy = e[0]; r->bond_energy = e[1];
        r->n_env = d->n_env;
        for (int k = 0; k < IGMK_MAX_ENVELOPES; k++) r->env_energy[k] = k < d->n_env ? e[2 + k] : 0.0;
        r->cg_iters = iter; r->cg_evals = neval; r->cg_stop = stop;
        r->n_rebuilds += c.n_rebuild;
        B.flags[blockIdx.x] |= (c.overflow ? 1 : 0) | (bad || !isfinite(ecur) ? 2 : 0);
    }
}

// ---------------------------------------------------------------------------------------
// M10: per-restraint violation ratios at the current coordinates, what ModelingStep.task gathers
// force by force (igm/steps/ModelingStep.py:345-372) through HarmonicUpperBound/LowerBound
// .getViolationRatio (igm/model/forces.py:128-136, :176-184) and EllipticEnvelope.getScores
// (:220-250).  Arithmetic as numpy does it on float32 positions: dist = np.linalg.norm of a float32
// 3-vector (float32 products, double accumulation, one rounding, float32 sqrt); then the float32
// chain score = k*(dist-d), ratio = score/(k*d) when d is a float32 value (Polymer, Hi-C, SPRITE:
// contactRange*(ri+rj) on float32 radii) or the same in float64 when it is not (FISH targets).
struct SubBond { int32_t i, j; double r0, k; int32_t style, f32; };

__device__ __forceinline__ float norm_f32_sdot(float dx, float dy, float dz)
{
    const double acc = __dadd_rn(__dadd_rn((double)__fmul_rn(dx, dx), (double)__fmul_rn(dy, dy)), (double)__fmul_rn(dz, dz));
    return __fsqrt_rn(__double2float_rn(acc));
}

__global__ void __launch_bounds__(256)
k_bond_violations(const float4* __restrict__ xyzr, const SubBond* __restrict__ bonds, const int32_t* __restrict__ atom_off_of_bond,
                  int64_t n_bonds, double* __restrict__ ratio)
{
    for (int64_t b = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; b < n_bonds; b += (int64_t)gridDim.x * blockDim.x) {
        const SubBond q = bonds[b];
        const float4 pi = xyzr[atom_off_of_bond[b] + q.i], pj = xyzr[atom_off_of_bond[b] + q.j];
        const float dist = norm_f32_sdot(__fsub_rn(pi.x, pj.x), __fsub_rn(pi.y, pj.y), __fsub_rn(pi.z, pj.z));
        double r = 0.0;
        if (q.r0 != 0.0) {                                                   // forces.py:136, :184
            if (q.f32) {
                const float d = (float)q.r0, k = (float)q.k;
                const bool viol = q.style == IGMK_BOND_UPPER ? !(dist <= d) : !(dist >= d);
                if (viol) {
                    const float diff = q.style == IGMK_BOND_UPPER ? __fsub_rn(dist, d) : __fsub_rn(d, dist);
                    r = (double)__fdiv_rn(__fmul_rn(k, diff), __fmul_rn(k, d));
                }
            } else {
                const double dd = (double)dist;
                const bool viol = q.style == IGMK_BOND_UPPER ? !(dd <= q.r0) : !(dd >= q.r0);
                if (viol) r = (q.k * (q.style == IGMK_BOND_UPPER ? dd - q.r0 : q.r0 - dd)) / (q.k * q.r0);
            }
        }
        ratio[b] = r;
    }
}

// Envelope members: u = (1 - 1/sqrt(q)) * |x| for k > 0 and q > 1, u = 1 - sqrt(q) for k < 0 and q < 1,
// else 0, with q = sqrt(sum x^2/(axis - r)^2) (forces.py:224-244; the reference takes the square root
// twice); the host finishes with the force's `scale`: ratio = max(0, t*k)/(k*scale), t = u/scale (k > 0).
struct EnvMember { int32_t atom; int32_t pad; double a, b, c, k; };

__global__ void __launch_bounds__(256)
k_env_violations(const float4* __restrict__ xyzr, const EnvMember* __restrict__ mem, int64_t n_mem, double* __restrict__ u_out)
{
    for (int64_t m = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; m < n_mem; m += (int64_t)gridDim.x * blockDim.x) {
        const EnvMember e = mem[m];
        const float4 p = xyzr[e.atom];
        const double r = (double)p.w;
        const double sa = e.a - r, sb = e.b - r, sc = e.c - r;
        const double s = __dadd_rn(__dadd_rn(__ddiv_rn((double)__fmul_rn(p.x, p.x), sa * sa), __ddiv_rn((double)__fmul_rn(p.y, p.y), sb * sb)),
                                   __ddiv_rn((double)__fmul_rn(p.z, p.z), sc * sc));
        const double q = sqrt(s);
        double u = 0.0;
        if (q > 1.0 && e.k > 0.0) u = (1.0 - 1.0 / sqrt(q)) * (double)norm_f32_sdot(p.x, p.y, p.z);
        else if (q < 1.0 && e.k < 0.0) u = 1.0 - sqrt(q);
        u_out[m] = u;
    }
}

}  // namespace igmk
