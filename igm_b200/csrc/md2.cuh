// md2.cuh -- the annealing-MD kernel of the M-step (lammps.py:285-351), one CTA per structure.
//
// Work decomposition: THREAD PER ATOM.  Thread t of 768 owns the atoms at positions t, t+768, ...
// (K slots per thread) of a host-chosen ordering in which the 32 atoms of a warp-slot own similar
// numbers of bonds.  Velocities and forces of a thread's atoms live in registers; shared memory holds
// what other threads read: the positions (float4 x,y,z,radius), the coordinates of the last neighbour
// build, the cell grid of that build and one 16-byte "inbox" of bond flags per atom.  Every
// floating-point sum is taken in a fixed order (no floating-point atomics): runs are bit-reproducible.
//
// The kernel is bound by instruction issue, so the force loops are organised around the COMMON case:
// in a population that is close to satisfying its restraints most pair candidates do not overlap and
// most bonds are not violated.  Every loop therefore first does the cheap part for four candidates
// (gather the partner position, squared distance, one comparison: ~11 instructions per candidate) and
// enters the expensive part (rsqrt / sin, force, accumulation) only when one of the four is active.
//
// Pairs (pair_style soft, lammps.py:189,216-219): neighbour table nl[batch][position][4] (uint16; 4
// neighbours per 8-byte load), all beads within cutoff + skin, rebuilt in-kernel through a cell grid in
// shared memory when some bead has moved half the skin.  Rows are written 8 bytes at a time.
//
// Bonds (harmonic_upper_bound / harmonic_lower_bound) are evaluated ONCE per bond: the host orients every
// bond -- one end OWNS it (walks it in its sliced-ELL row, ell[warp-slot][batch][lane][4], one 16-byte load
// per 4 items), the other end RECEIVES it through slot q < 15 of its inbox.  When the owner finds the
// restraint violated it adds the force to its own accumulator, writes (dx, dy, dz, f/r) to the mailbox
// gf[position*16 + q] in global memory and sets flag byte q of the receiver's inbox.  After one barrier every
// atom reads its 16 flag bytes with one LDS.128, collects the flagged mailboxes in slot order (a
// deterministic sum) and clears the inbox.  The receiving side thus costs a few instructions per ATOM and
// evaluation instead of a full evaluation per bond.  Bonds that do not fit an inbox (more than 15 received)
// are listed at both ends as before (SYM items).  Items of the most frequent bond type (an upper bound: all
// polymer and Hi-C restraints of a uniform-radius genome) sit in block A of a warp-slot and run a loop
// specialised for that type; everything else is in block B.  A bound that is inactive with a margin of one
// skin cannot switch on before the next rebuild, so the first evaluation after a rebuild walks the full
// lists and writes the candidates to `eact`, which the following evaluations walk.
#pragma once
#include <type_traits>
#include "mstep.cuh"

namespace igmk {

#define MOVES(k) ((meta[k] & 3u) == IGMK_KIND_BEAD || (meta[k] & 3u) == IGMK_KIND_CENTROID)

#ifndef MD2_THREADS
#define MD2_THREADS 1024
#endif
constexpr int MD2_T = MD2_THREADS;    // threads per CTA of the cluster and global-memory kernels (md_cluster.cuh, md_big.cuh)
constexpr int MD2_NW = MD2_T / 32;    // warps per CTA
// k_md2 runs 768 threads: 80 registers per thread instead of 64 (no spills in the force loops, loads further ahead of
// their use) are worth more than a fourth more warps.  Measured on the 3008-bead bench structures, same positions per CTA
// (3072): 512 x 6 / 640 x 5 / 768 x 4 / 1024 x 3 threads x atoms per thread -> 68.5 / 67.0 / 66.7 / 71.9 ms
// (tools/prof_md.py 148 0.06); 896 x 4 = 3584 positions: 76.0.
#ifndef MD2S_THREADS
#define MD2S_THREADS 768
#endif
constexpr int MD2S_T = MD2S_THREADS;
constexpr int MD2S_NW = MD2S_T / 32;
constexpr int MD2_KMAX = 5;           // slots per thread supported (n_atoms <= 3840 on the 1-CTA path: shared memory)
constexpr int MD2_NLW0 = 256;         // initial neighbour-table width (columns)
constexpr uint16_t MD2_NONE = 0xffffu;
constexpr int MD2_INBOX = 16;         // flag bytes per atom; slot 15 is never used (padding items point at it)
// bond item of k_md2: bits 0-3 inbox slot q, 4-16 partner position, 17 SYM (listed at both ends, no inbox), 18-31 type
constexpr uint32_t MD2_IT_ADDR = 0x1ffffu, MD2_IT_POS = 0x1fff0u, MD2_IT_SYM = 0x20000u;
constexpr int MD2_IT_TSHIFT = 18;
constexpr int MD2_MAX_TYPES = 1 << 14;

struct Md2View {
    const uint16_t* perm;      // [sum np] position -> original atom (0xffff: empty position)
    const int32_t*  ell_off;   // k_md_big: per structure 32*K+1 offsets (in items) of each warp-slot block
    const uint32_t* ell_items; // sliced-ELL bond items
    uint32_t*       ell_act;   // same size: the items that can be active before the next rebuild (kernel scratch)
    uint16_t*       nl;        // neighbour table [sum np * nl_width], column-major per structure
    int32_t         nl_width;
    long long*      prof;      // optional [n_struct][16] SM cycles of thread 0 (IGMK_PROF=1)
    // k_md2 only
    const int4*     eblk;      // per structure K*32 entries {offset A, batches A, offset B, batches B} (offsets in items)
    float4*         gf;        // bond mailboxes [sum np * 16]
    double*         dbg_forces;// debug: [atoms*3] forces of the run-setup evaluation of stage MdParams.debug_stage
};

struct Md2Desc {               // per structure
    int64_t ell_item_off;
    int32_t perm_off;          // offset into perm (== sum of np of the structures before)
    int32_t ell_desc_off;      // offset into ell_off (k_md_big) / eblk (k_md2)
    int32_t K;
    int32_t pad;
};

// single-instruction approximations (MUFU) without the denormal fix-up code of the CUDA intrinsics;
// their ~1e-6 relative error is below the float32 round-off the MD trajectory carries anyway
__device__ __forceinline__ float fast_rsqrt(float v) { float r; asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(v)); return r; }
__device__ __forceinline__ float fast_rcp(float v) { float r; asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(v)); return r; }
__device__ __forceinline__ float fast_sin(float v) { float r; asm("sin.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(v)); return r; }
// shared-memory load through an explicit 32-bit shared address (keeps the address arithmetic to one op)
__device__ __forceinline__ float4 lds_f4(uint32_t addr)
{
    float4 v;
    asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(addr));
    return v;
}

__host__ __device__ inline size_t md2_smem_bytes(int K)
{
    const size_t np = (size_t)K * MD2S_T;
    size_t b = (np + 1) * 16;                 // x + one sentinel position far away (padding of the neighbour rows)
    b += np * 16;                             // positions in cell order (neighbour search) / bond inboxes (between rebuilds)
    b += np * 12;                             // build coordinates
    b += ((size_t)MAX_CELLS + 4) * 4;         // cell_start (+1 entry, padded to 16 bytes)
    b += np * 2 * 2;                          // cell_atoms, cell_of (= row lengths between rebuilds)
    b += 2 * 32 * 8 * 4;                      // reduction scratch
    b += IGMK_MAX_ENVELOPES * 4 * 4;          // envelope parameters
    b += (size_t)K * MD2S_NW * 16;             // bond block descriptors
    b += (size_t)K * MD2S_NW * 8;              // candidate batches per warp-slot (A, B)
    b += 16 * 16;                             // compaction table of the neighbour search
    return (b + 255) / 256 * 256;
}

// Neighbour row of one atom, written 4 entries (one 8-byte store) at a time: the table is [batch][position][4] uint16, so
// the 8-byte stores of neighbouring positions share sectors -- 2-byte stores, one sector each, saturate the L2 instead.
struct Md2Row {
    uint2* base; size_t stride; uint32_t lo, hi; int cnt, cap;
    __device__ __forceinline__ Md2Row(uint16_t* table, int p, int np, int width)
        : base((uint2*)table + p), stride((size_t)np), lo(0u), hi(0u), cnt(0), cap(width) {}
    __device__ __forceinline__ void add(uint32_t v)
    {
        const int slot = cnt & 3;
        const uint32_t sv = v << ((slot & 1) << 4);
        if (slot < 2) lo |= sv; else hi |= sv;
        if (slot == 3) {
            if (cnt < cap) base[(size_t)(cnt >> 2) * stride] = make_uint2(lo, hi);
            lo = 0u; hi = 0u;
        }
        cnt++;
    }
    // pad the last batch with `self` (an index that never interacts); -> entries kept, *ovf set when the row did not fit
    __device__ __forceinline__ int finish(uint32_t self, int* ovf)
    {
        const int n = cnt;
        while (cnt & 3) add(self);
        if (n > cap) { *ovf = 1; return cap; }
        return n;
    }
};

// The same row filled up to four entries at a time from the neighbour search: `packed` holds n (0..4) indices in its low
// 16-bit lanes, the other lanes zero.  No loop over the hits, no branch but the store.
struct Md2RowPacked {
    uint2* base; size_t stride; uint64_t acc; int fill, batches, cap4;
    __device__ __forceinline__ Md2RowPacked(uint16_t* table, int p, int np, int width)
        : base((uint2*)table + p), stride((size_t)np), acc(0ull), fill(0), batches(0), cap4(width >> 2) {}
    __device__ __forceinline__ void append(uint64_t packed, int n)
    {
        const uint64_t comb = acc | (packed << (fill << 4));
        const int tot = fill + n;
        if (tot >= 4) {
            if (batches < cap4) base[(size_t)batches * stride] = make_uint2((uint32_t)comb, (uint32_t)(comb >> 32));
            batches++;
            acc = (packed >> ((3 - fill) << 4)) >> 16;       // what did not fit (two shifts: fill may be 0)
            fill = tot - 4;
        } else { acc = comb; fill = tot; }
    }
    // pad the last batch with `self`; -> entries kept, *ovf set when the row did not fit
    __device__ __forceinline__ int finish(uint32_t self, int* ovf)
    {
        const int n = (batches << 2) + fill;
        if (fill > 0) {
            const uint64_t s4 = (uint64_t)self * 0x0001000100010001ull;
            const uint64_t comb = acc | (s4 << (fill << 4));
            if (batches < cap4) base[(size_t)batches * stride] = make_uint2((uint32_t)comb, (uint32_t)(comb >> 32));
        }
        if (n > (cap4 << 2)) { *ovf = 1; return cap4 << 2; }
        return n;
    }
};

// PROF: per-phase clock64 counters (IGMK_PROF=1); a separate instantiation, because the counters cost ~14 registers
template <int K, bool PROF = false>
__global__ void __launch_bounds__(MD2S_T, 1)
k_md2(Batch B, Md2View V, const Md2Desc* __restrict__ mdesc, MdParams P)
{
    extern __shared__ __align__(16) unsigned char smem[];
    constexpr int np = K * MD2S_T;
    constexpr int CELL_CAP = (np * 4 < MAX_CELLS) ? np * 4 : MAX_CELLS;   // the cell cursors of the rebuild live in the cell-ordered copy (16 B per position)
    const StructDesc* d = B.desc + blockIdx.x;
    const Md2Desc md = mdesc[blockIdx.x];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int n = d->n_atoms;

    float4* x = (float4*)smem;
    float4* xsort = x + np + 1;               // rebuild only (x[np] is the far-away sentinel the neighbour rows are padded with)
    uint4* inbox = (uint4*)xsort;             // [np] 16 flag bytes per atom, between rebuilds
    float* xbx = (float*)(xsort + np); float* xby = xbx + np; float* xbz = xby + np;
    uint32_t* cell_start = (uint32_t*)(xbz + np);
    uint16_t* cell_atoms = (uint16_t*)(cell_start + MAX_CELLS + 4);
    uint16_t* cell_of = cell_atoms + np;      // during a rebuild: cell of a position; afterwards: its row length
    float* rbuf = (float*)(cell_of + np);
    float* envp = rbuf + 2 * 256;
    int4* eblk = (int4*)(envp + IGMK_MAX_ENVELOPES * 4);    // [K*32]
    int2* bact = (int2*)(eblk + K * MD2S_NW);                 // [K*32] candidate batches (A, B)
    uint4* lut = (uint4*)(bact + K * MD2S_NW);                // [16] hit mask of four candidates -> byte selectors that pack the hits
    unsigned char* flagb = (unsigned char*)inbox;
    int rb_par = 0;

    const igmk_protocol_t& pr = P.prot;
    const uint16_t* perm = V.perm + md.perm_off;
    uint32_t* eact = V.ell_act + md.ell_item_off;
    const uint32_t* eitems = V.ell_items + md.ell_item_off;
    uint16_t* nl = V.nl + (size_t)md.perm_off * V.nl_width;
    const int nlw = V.nl_width;
    float4* gf = V.gf + (size_t)md.perm_off * MD2_INBOX;
    const float2* btype = B.btype_f + d->btype_off;
    float4* xg = B.xyzr + d->atom_off;

    // ---- per-slot state in registers ----
    uint32_t meta[K];         // bits 0-1: IGMK_KIND_*, 3 = empty position; bits 8-15: envelope membership mask
    float vx[K], vy[K], vz[K], fx[K], fy[K], fz[K];
#pragma unroll
    for (int k = 0; k < K; k++) {
        const int p = k * MD2S_T + tid;
        const uint16_t a = perm[p];
        float4 q = make_float4(0.f, 0.f, 0.f, 0.f);
        meta[k] = 3u;
        if (a != MD2_NONE) {
            q = xg[a];
            meta[k] = (uint32_t)B.kind[d->atom_off + a] | (B.envmask[d->atom_off + a] << 8);
        }
        x[p] = q;
        vx[k] = vy[k] = vz[k] = 0.f; fx[k] = fy[k] = fz[k] = 0.f;
    }
    int uni = 1;
#pragma unroll
    for (int k = 0; k < K; k++) {
        const int p = k * MD2S_T + tid;
        if ((meta[k] & 3u) == IGMK_KIND_BEAD && x[p].w != d->maxrad) uni = 0;
    }
    if (tid < K * MD2S_NW) { eblk[tid] = V.eblk[md.ell_desc_off + tid]; bact[tid] = make_int2(0, 0); }
    if (tid == 0) x[np] = make_float4(1e18f, 1e18f, 1e18f, 0.f);
    if (tid < 16) {          // candidates j = 0..3 sit in 16-bit lane j of (i01, i23); lane t of the output takes the t-th set bit
        uint32_t sel[4] = {0u, 0u, 0u, 0u};
        int n = 0;
        for (int j = 0; j < 4; j++) if (tid & (1 << j)) sel[n++] = (uint32_t)(2 * j) | ((uint32_t)(2 * j + 1) << 4);
        lut[tid] = make_uint4(sel[0] | (sel[1] << 8), sel[2] | (sel[3] << 8),
                              n >= 2 ? 0xffffffffu : (n == 1 ? 0xffffu : 0u), n >= 4 ? 0xffffffffu : (n == 3 ? 0xffffu : 0u));
    }
    const float2 bt0 = btype[0];       // block A's type: the most frequent one (the host orders the table by use)
    const float skin = d->maxrad * P.skin_factor;
    const float lim = 0.25f * skin * skin;
    const float bond_pref2 = 2.f * (float)pr.bond_prefactor;
    float ev_c = -1.f;                 // evfactor * stage scale / pi  (force prefactor / rc), < 0: no sterics
    int n_rebuild = 0, overflow = 0;
    bool bond_build = true;
    const bool uniform_rc = __syncthreads_and(uni) != 0 && !P.general_pairs;
    long long t_search = 0, t_recv = 0, t_bondA = 0, t_pair = 0, t_force = 0, t_build = 0, t_smax = 0, t_ssum = 0, t_fmax = 0, t_fsum = 0, n_cand = 0, n_full = 0, n_nbr = 0;
    auto tick = [&]() -> long long { return PROF ? clock64() : 0ll; };
    __syncthreads();

    // ------------------------------------------------------------------------------------
    auto block_sum1 = [&](float v) -> float {
        float* buf = rbuf + (rb_par ^= 1) * 256;
        v = warp_sum(v);
        if (lane == 0) buf[warp] = v;
        __syncthreads();
        return warp_sum(lane < MD2S_NW ? buf[lane] : 0.f);     // <= 32 partials: one per lane, same tree in every warp
    };

    auto set_stage = [&](double ev_scale, double env_scale) {
        ev_c = d->evfactor < 0 ? -1.f : (float)(d->evfactor * ev_scale / 3.14159265358979323846);
        if (tid < IGMK_MAX_ENVELOPES * 4) {
            const int e = tid >> 2, c = tid & 3;
            envp[tid] = e < d->n_env ? (float)(c < 3 ? d->env[e][c] * env_scale : d->env[e][3]) : 0.f;
        }
        __syncthreads();
    };

    const uint32_t xs = (uint32_t)__cvta_generic_to_shared(x);

    // ---- neighbour table: cell binning in shared memory, cutoff + skin; resets the bond inboxes ----
    auto rebuild = [&]() {
        const long long tb0 = tick();
        float lo[3] = {1e30f, 1e30f, 1e30f}, hi[3] = {-1e30f, -1e30f, -1e30f};
#pragma unroll
        for (int k = 0; k < K; k++) {
            const int p = k * MD2S_T + tid;
            const float4 q = x[p];
            if ((meta[k] & 3u) == IGMK_KIND_BEAD) {
                lo[0] = fminf(lo[0], q.x); hi[0] = fmaxf(hi[0], q.x);
                lo[1] = fminf(lo[1], q.y); hi[1] = fmaxf(hi[1], q.y);
                lo[2] = fminf(lo[2], q.z); hi[2] = fmaxf(hi[2], q.z);
            }
        }
        float* rb = rbuf + (rb_par ^= 1) * 256;
#pragma unroll
        for (int c = 0; c < 3; c++) { lo[c] = warp_min(lo[c]); hi[c] = warp_max(hi[c]); }
        if (lane == 0) {
#pragma unroll
            for (int c = 0; c < 3; c++) { rb[warp * 6 + c] = lo[c]; rb[warp * 6 + 3 + c] = hi[c]; }
        }
        for (int c = tid; c <= MAX_CELLS; c += MD2S_T) cell_start[c] = 0;
        __syncthreads();
#pragma unroll
        for (int c = 0; c < 3; c++) {
            lo[c] = warp_min(lane < MD2S_NW ? rb[lane * 6 + c] : 1e30f); hi[c] = warp_max(lane < MD2S_NW ? rb[lane * 6 + 3 + c] : -1e30f);
        }
        const float rlist = 2.f * d->maxrad + skin, rlist2 = rlist * rlist;
        int dim[3]; float inv[3];
#pragma unroll
        for (int c = 0; c < 3; c++) {
            float ext = hi[c] - lo[c];
            if (!(ext > 0.f)) ext = 0.f;
            int m = (int)(ext / rlist) + 1;
            if (m > MAX_CELL_DIM) m = MAX_CELL_DIM;
            float cw = ext / (float)m;
            if (cw < rlist) cw = rlist;
            dim[c] = m; inv[c] = 1.f / cw;
        }
        while (dim[0] * dim[1] * dim[2] > CELL_CAP) {            // (one-slot structures only: fewer, wider cells)
            const int c = dim[0] >= dim[1] && dim[0] >= dim[2] ? 0 : (dim[1] >= dim[2] ? 1 : 2);
            dim[c]--;
            float ext = hi[c] - lo[c];
            if (!(ext > 0.f)) ext = 0.f;
            float cw = ext / (float)dim[c];
            if (cw < rlist) cw = rlist;
            inv[c] = 1.f / cw;
        }
        const int ncell = dim[0] * dim[1] * dim[2];
        // counts per cell (integer atomics: the counts do not depend on the order)
#pragma unroll
        for (int k = 0; k < K; k++) {
            const int p = k * MD2S_T + tid;
            int c = MD2_NONE;
            if ((meta[k] & 3u) == IGMK_KIND_BEAD) {
                const float4 q = x[p];
                int cx = (int)((q.x - lo[0]) * inv[0]), cy = (int)((q.y - lo[1]) * inv[1]), cz = (int)((q.z - lo[2]) * inv[2]);
                cx = min(max(cx, 0), dim[0] - 1); cy = min(max(cy, 0), dim[1] - 1); cz = min(max(cz, 0), dim[2] - 1);
                c = (cz * dim[1] + cy) * dim[0] + cx;
                atomicAdd(&cell_start[c + 1], 1u);
            }
            cell_of[p] = (uint16_t)c;
        }
        __syncthreads();
        // inclusive scan of cell_start[1..ncell] (cell_start[c+1] held the count of cell c)
        {
            const int per = (ncell + MD2S_T - 1) / MD2S_T;
            const int c0 = min(tid * per, ncell), c1 = min(c0 + per, ncell);
            int s = 0;
            for (int c = c0; c < c1; c++) s += cell_start[c + 1];
            int incl = s;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) { int t = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += t; }
            int* ib = (int*)(rbuf + (rb_par ^= 1) * 256);
            if (lane == 31) ib[warp] = incl;
            __syncthreads();
            int base = 0;
            for (int w = 0; w < warp; w++) base += ib[w];
            int run = base + incl - s;
            for (int c = c0; c < c1; c++) { const int cnt = cell_start[c + 1]; run += cnt; cell_start[c + 1] = run; }
        }
        __syncthreads();
        // drop atoms unordered behind the cell start, then sort each (tiny) cell by index so that the
        // row order, and with it the floating-point summation order, is independent of the race.
        uint32_t* cursor = (uint32_t*)xsort;        // np * 16 bytes >= 4 * MAX_CELLS for every K; the cell-ordered copy is written later
        for (int c = tid; c < ncell; c += MD2S_T) cursor[c] = 0;
        __syncthreads();
#pragma unroll
        for (int k = 0; k < K; k++) {
            const int p = k * MD2S_T + tid;
            const int c = cell_of[p];
            if (c != MD2_NONE) {
                const uint32_t pos = atomicAdd(&cursor[c], 1u);
                cell_atoms[cell_start[c] + pos] = (uint16_t)p;
            }
        }
        __syncthreads();
        for (int c = tid; c < ncell; c += MD2S_T) {
            const int b = cell_start[c], e = cell_start[c + 1];
            for (int q = b + 1; q < e; q++) {
                const uint16_t v = cell_atoms[q]; int r = q - 1;
                while (r >= b && cell_atoms[r] > v) { cell_atoms[r + 1] = cell_atoms[r]; r--; }
                cell_atoms[r + 1] = v;
            }
        }
        __syncthreads();
#pragma unroll
        for (int k = 0; k < K; k++) {
            const int p = k * MD2S_T + tid;
            const float4 q = x[p];
            xbx[p] = q.x; xby[p] = q.y; xbz[p] = q.z;
        }
        // positions in cell order: the search below walks them contiguously
        const long long ts0 = tick();
        const int nbead = cell_start[ncell];
        for (int s0 = tid; s0 < nbead; s0 += MD2S_T) {      // (the radius is not needed here: w carries the atom's position index)
            const uint32_t a = cell_atoms[s0];
            float4 v = x[a];
            v.w = __uint_as_float(a);
            xsort[s0] = v;
        }
        __syncthreads();
        // Thread t takes the t-th bead IN CELL ORDER (not one of its own atoms), so the lanes of a warp sit in
        // the same or adjacent cells, scan nearly the same ranges (little divergence, broadcast shared-memory
        // reads) and need no index indirection until a candidate is accepted.
        int ovf = 0;
        const uint32_t xsa = (uint32_t)__cvta_generic_to_shared(xsort);
        for (int s0 = tid; s0 < nbead; s0 += MD2S_T) {
            const int p = cell_atoms[s0];
            const float4 q = xsort[s0];
            const int c = cell_of[p];
            const int cx = c % dim[0], cy = (c / dim[0]) % dim[1], cz = c / (dim[0] * dim[1]);
            const int ax0 = max(cx - 1, 0), ax1 = min(cx + 1, dim[0] - 1);
            Md2RowPacked row(nl, p, np, nlw);
            for (int dz = -1; dz <= 1; dz++) {
                const int az = cz + dz;
                if (az < 0 || az >= dim[2]) continue;
                for (int dy = -1; dy <= 1; dy++) {
                    const int ay = cy + dy;
                    if (ay < 0 || ay >= dim[1]) continue;
                    const int cb = (az * dim[1] + ay) * dim[0];
                    const int b = cell_start[cb + ax0], e = cell_start[cb + ax1 + 1];   // 3 cells, contiguous
                    for (int s1 = b; s1 < e; s1 += 4) {
                        // four candidates at once, branch-free: the four loads go out together (a test that short-circuits on
                        // s2 < e compiles to a branch around each load and serialises the four shared-memory latencies)
                        float4 pj[4];
#pragma unroll
                        for (int u = 0; u < 4; u++) pj[u] = lds_f4(xsa + ((uint32_t)(s1 + u) << 4));   // may run past e: still inside this CTA's shared memory, masked below
                        uint32_t hits = 0u;
#pragma unroll
                        for (int u = 0; u < 4; u++) {
                            const float ddx = q.x - pj[u].x, ddy = q.y - pj[u].y, ddz = q.z - pj[u].z;
                            // list radius of the largest beads: a superset of cutoff + skin of this pair when radii differ
                            hits |= (uint32_t)(ddx * ddx + ddy * ddy + ddz * ddz < rlist2) << u;
                        }
                        const int rem = e - s1;
                        const uint32_t me = (uint32_t)(s0 - s1);
                        hits &= (rem >= 4 ? 0xfu : (1u << rem) - 1u) & (me < 4u ? ~(1u << me) : ~0u);   // past the range; myself
                        // the hits' indices (the w of the cell-ordered copy), packed in candidate order through the table
                        const uint32_t i01 = __byte_perm(__float_as_uint(pj[0].w), __float_as_uint(pj[1].w), 0x5410);
                        const uint32_t i23 = __byte_perm(__float_as_uint(pj[2].w), __float_as_uint(pj[3].w), 0x5410);
                        const uint4 L = lut[hits];
                        const uint32_t r0 = __byte_perm(i01, i23, L.x) & L.z, r1 = __byte_perm(i01, i23, L.y) & L.w;
                        row.append(((uint64_t)r1 << 32) | r0, __popc(hits));
                    }
                }
            }
            cell_of[p] = (uint16_t)row.finish((uint32_t)np, &ovf);  // padded with the sentinel; the row length stays here until the next rebuild
        }
        if (PROF) {                                   // how unevenly the search is spread over the warps
            if (lane == 0) ((long long*)rbuf)[warp] = tick() - ts0;
            __syncthreads();
            if (tid == 0) { long long mx = 0, sm = 0; for (int q = 0; q < MD2S_NW; q++) { const long long v = ((long long*)rbuf)[q]; mx = v > mx ? v : mx; sm += v; } t_smax += mx; t_ssum += sm / MD2S_NW; }
        }
        __syncthreads();
        // the cell-ordered copy is dead: its memory becomes the bond inboxes (all flags clear); dummies / empty positions: no row
#pragma unroll
        for (int k = 0; k < K; k++) {
            const int p = k * MD2S_T + tid;
            inbox[p] = make_uint4(0u, 0u, 0u, 0u);
            if (cell_of[p] == MD2_NONE) cell_of[p] = 0;
        }
        overflow |= __syncthreads_or(ovf);
        if (PROF) { t_search += tick() - ts0; t_build += tick() - tb0; }
        bond_build = true;        // the next force evaluation also refreshes the bond candidate lists
        n_rebuild++;
    };

    // ---- forces on my atoms ---------------------------------------------------------------
    // Rows of the neighbour table are padded to a multiple of 4 with the index of a sentinel position far away and
    // the bond blocks with items to oneself (r2 = 0, never beyond an upper bound): the loops need no bounds tests.
    // BUILD: first evaluation after a rebuild -- walk the full bond lists and write the candidate lists.
    auto forces = [&](auto build_tag) {
        constexpr bool BUILD = decltype(build_tag)::value;
        const long long tf0 = tick();
#pragma unroll
        for (int k = 0; k < K; k++) {
            const int p = k * MD2S_T + tid;
            const float4 pi = x[p];
            float ax = 0.f, ay = 0.f, az = 0.f;
            const long long tp1 = tick();
            // harmonic upper / lower bounds I own: F = -2 pref k (r - r0) r_hat when violated
            const int g = k * MD2S_NW + warp;
            const int4 eb = eblk[g];
            const int2 ba = bact[g];
            // first batch of my neighbour row: in flight while the bond loops run
            const uint2* col = (const uint2*)nl + p;
            const int nbp = ev_c >= 0.f ? ((int)cell_of[p] + 3) >> 2 : 0;
            uint2 wn = nbp > 0 ? col[0] : make_uint2(0u, 0u);
            {   // block A: every item is of type 0, an upper bound, and has an inbox slot at the other end
                const uint4* it = (const uint4*)((BUILD ? eitems : eact) + eb.x) + lane;   // [batch][lane][4 items]
                const int nb4 = BUILD ? eb.y : ba.x;
                if (PROF) { if (BUILD) n_full += nb4; else n_cand += nb4; n_nbr += nbp; }
                uint32_t* out = eact + eb.x + (lane << 2);
                const float r0 = bt0.x, r02 = r0 * r0, kf = -bond_pref2 * bt0.y;
                const float cl = fmaxf(r0 - skin, 0.f), cl2 = cl * cl;
                int ccnt = 0;
                uint4 nxt = nb4 > 0 ? it[0] : make_uint4(0, 0, 0, 0);
                for (int c = 0; c < nb4; c++) {
                    const uint4 cur = nxt;
                    if (c + 1 < nb4) nxt = it[(c + 1) * 32];                  // next batch in flight while this one is used
                    const uint32_t v[4] = {cur.x, cur.y, cur.z, cur.w};
                    float4 pj[4];
#pragma unroll
                    for (int u = 0; u < 4; u++) pj[u] = lds_f4(xs + (v[u] & MD2_IT_POS));
                    float dx[4], dy[4], dz[4], r2[4];
                    bool any = false;
#pragma unroll
                    for (int u = 0; u < 4; u++) {
                        dx[u] = pi.x - pj[u].x; dy[u] = pi.y - pj[u].y; dz[u] = pi.z - pj[u].z;
                        r2[u] = dx[u] * dx[u] + dy[u] * dy[u] + dz[u] * dz[u];
                        any |= r2[u] > r02;                                  // a self item has r2 = 0
                    }
                    if (any) {                               // a restraint of the four is violated
#pragma unroll
                        for (int u = 0; u < 4; u++) {
                            if (r2[u] > r02) {
                                const float fb = kf * (1.f - r0 * fast_rsqrt(r2[u]));
                                ax += dx[u] * fb; ay += dy[u] * fb; az += dz[u] * fb;
                                gf[v[u] & MD2_IT_ADDR] = make_float4(dx[u], dy[u], dz[u], fb);
                                flagb[v[u] & MD2_IT_ADDR] = 1;
                            }
                        }
                    }
                    if (BUILD) {
#pragma unroll
                        for (int u = 0; u < 4; u++)
                            if (r2[u] > cl2) { out[((ccnt >> 2) << 7) + (ccnt & 3)] = v[u]; ccnt++; }
                    }
                }
                if (BUILD) {
                    const int most = warp_max((ccnt + 3) >> 2);
                    for (int cc = ccnt; cc < (most << 2); cc++) out[((cc >> 2) << 7) + (cc & 3)] = ((uint32_t)p << 4) | 15u;   // self item
                    if (lane == 0) bact[g].x = most;
                }
            }
            if (PROF) t_bondA += tick() - tp1;
            if (eb.w > 0) {   // block B: any type, upper or lower bound; SYM items are evaluated at both ends
                const uint4* it = (const uint4*)((BUILD ? eitems : eact) + eb.z) + lane;
                const int nb4 = BUILD ? eb.w : ba.y;
                uint32_t* out = eact + eb.z + (lane << 2);
                int ccnt = 0;
                for (int c = 0; c < nb4; c++) {
                    const uint4 cur = it[c * 32];
                    const uint32_t v[4] = {cur.x, cur.y, cur.z, cur.w};
                    float4 pj[4];
#pragma unroll
                    for (int u = 0; u < 4; u++) pj[u] = lds_f4(xs + (v[u] & MD2_IT_POS));
#pragma unroll
                    for (int u = 0; u < 4; u++) {
                        const float2 t = btype[v[u] >> MD2_IT_TSHIFT];     // (r0, k); k < 0: lower bound
                        const float dx = pi.x - pj[u].x, dy = pi.y - pj[u].y, dz = pi.z - pj[u].z;
                        const float r2 = dx * dx + dy * dy + dz * dz;
                        const float r02 = t.x * t.x;
                        if ((t.y > 0.f ? (r2 > r02) : (r2 < r02)) && r2 > 0.f) {
                            const float fb = -bond_pref2 * fabsf(t.y) * (1.f - t.x * fast_rsqrt(r2));
                            ax += dx * fb; ay += dy * fb; az += dz * fb;
                            if (!(v[u] & MD2_IT_SYM)) {
                                gf[v[u] & MD2_IT_ADDR] = make_float4(dx, dy, dz, fb);
                                flagb[v[u] & MD2_IT_ADDR] = 1;
                            }
                        }
                        if (BUILD) {
                            const float lo = fmaxf(t.x - skin, 0.f), hi = t.x + skin;
                            if (t.y > 0.f ? (r2 > lo * lo) : (r2 < hi * hi)) { out[((ccnt >> 2) << 7) + (ccnt & 3)] = v[u]; ccnt++; }
                        }
                    }
                }
                if (BUILD) {
                    const int most = warp_max((ccnt + 3) >> 2);
                    for (int cc = ccnt; cc < (most << 2); cc++) out[((cc >> 2) << 7) + (cc & 3)] = ((uint32_t)p << 4) | MD2_IT_SYM;   // self item
                    if (lane == 0) bact[g].y = most;
                }
            }
            const long long tp0 = tick();     // (the pair loop runs after the bond loops: the mailbox stores land meanwhile)
            // pair_style soft: F = A pi/rc sin(pi r/rc) r_hat, A = rc^2 ev/pi^2  =>  F/r = ev_c rc sin(pi r/rc) / r.  Branch-free:
            // at MD temperatures nearly every batch of four has an overlapping bead in some lane of the warp, so an early-out
            // only adds divergence here (measured, profiles/r02g: +8 % instructions).
            if (uniform_rc) {
                // every bead has the same radius (a genome binned at one resolution): the contact distance and with it
                // pi / rc and ev rc are constants of the loop -- 27 instead of 31 instructions per candidate (SASS), one MUFU fewer
                const float rcu = 2.f * d->maxrad, rc2 = rcu * rcu, a1 = 3.14159265358979323846f / rcu, a2 = ev_c * rcu;
                for (int q = 0; q < nbp; q++) {
                    const uint2 w = wn;
                    if (q + 1 < nbp) wn = col[(size_t)(q + 1) * np];            // next batch in flight while this one is used
                    const uint32_t j[4] = {w.x & 0xffffu, w.x >> 16, w.y & 0xffffu, w.y >> 16};
#pragma unroll
                    for (int u = 0; u < 4; u++) {
                        const float4 pj = lds_f4(xs + (j[u] << 4));
                        const float dx = pi.x - pj.x, dy = pi.y - pj.y, dz = pi.z - pj.z;
                        const float r2 = dx * dx + dy * dy + dz * dz;
                        const float rinv = fast_rsqrt(r2);
                        float fp = a2 * (fast_sin((r2 * rinv) * a1) * rinv);
                        fp = (r2 < rc2 && r2 > 0.f) ? fp : 0.f;                   // (the padding sentinel is far away)
                        ax += dx * fp; ay += dy * fp; az += dz * fp;
                    }
                }
            } else {
                for (int q = 0; q < nbp; q++) {
                    const uint2 w = wn;
                    if (q + 1 < nbp) wn = col[(size_t)(q + 1) * np];
                    const uint32_t j[4] = {w.x & 0xffffu, w.x >> 16, w.y & 0xffffu, w.y >> 16};
#pragma unroll
                    for (int u = 0; u < 4; u++) {
                        const float4 pj = lds_f4(xs + (j[u] << 4));
                        const float dx = pi.x - pj.x, dy = pi.y - pj.y, dz = pi.z - pj.z;
                        const float rc = pi.w + pj.w;
                        const float r2 = dx * dx + dy * dy + dz * dz;
                        const float rinv = fast_rsqrt(r2);
                        const float sn = fast_sin(3.14159265358979323846f * (r2 * rinv) * fast_rcp(rc));
                        float fp = ev_c * rc * sn * rinv;
                        fp = (r2 < rc * rc && r2 > 0.f) ? fp : 0.f;
                        ax += dx * fp; ay += dy * fp; az += dz * fp;
                    }
                }
            }
            if (PROF) t_pair += tick() - tp0;
            // fix ellipsoidalenvelope
            uint32_t m = meta[k] >> 8;
            while (m) {
                const int e = __ffs(m) - 1; m &= m - 1;
                const float ia = fast_rcp(envp[4 * e] - pi.w), ib = fast_rcp(envp[4 * e + 1] - pi.w),
                            ic = fast_rcp(envp[4 * e + 2] - pi.w), kk = envp[4 * e + 3];
                const float ux = pi.x * ia, uy = pi.y * ib, uz = pi.z * ic;
                const float k2 = ux * ux + uy * uy + uz * uz;
                const bool act = (kk > 0.f ? (k2 > 1.f) : (kk < 0.f && k2 < 1.f)) && k2 > 0.f;
                if (act) {
                    const float g2 = fabsf(kk) * (1.f - fast_rsqrt(k2));
                    ax -= g2 * pi.x; ay -= g2 * pi.y; az -= g2 * pi.z;
                }
            }
            fx[k] = ax; fy[k] = ay; fz[k] = az;
        }
        bond_build = false;
        const long long tr0 = tick();
        if (PROF) {                                   // how unevenly the force loops are spread over the warps
            if (lane == 0) ((long long*)rbuf)[64 + warp] = tr0 - tf0;
            __syncthreads();
            if (tid == 0) { long long mx = 0, sm = 0; for (int q = 0; q < MD2S_NW; q++) { const long long v = ((long long*)rbuf)[64 + q]; mx = v > mx ? v : mx; sm += v; } t_fmax += mx; t_fsum += sm / MD2S_NW; }
        }
        __syncthreads();            // every inbox flag and mailbox of this evaluation is written
        // bonds I receive: the owner's (dx, dy, dz, f/r) with the opposite sign, in inbox order; then clear the inbox.
        // Flag bytes (0 / 1) -> one bit per slot.  The first flagged mailbox of EVERY slot of this thread is requested before
        // any is used (the K round trips to L2 overlap); what is left of an atom's flags (few atoms have more than one
        // violated incoming bond at a time) follows four at a time in slot order.
        uint32_t mk[K];
        float4 f0[K];
#pragma unroll
        for (int k = 0; k < K; k++) {
            const int p = k * MD2S_T + tid;
            const uint4 fl = inbox[p];
            mk[k] = (((fl.x * 0x01020408u) >> 24) & 0xfu) | ((((fl.y * 0x01020408u) >> 24) & 0xfu) << 4) |
                    ((((fl.z * 0x01020408u) >> 24) & 0xfu) << 8) | ((((fl.w * 0x01020408u) >> 24) & 0xfu) << 12);
            f0[k] = make_float4(0.f, 0.f, 0.f, 0.f);
            if (mk[k]) {
                f0[k] = __ldcg(gf + ((size_t)p << 4) + (__ffs(mk[k]) - 1));
                mk[k] &= mk[k] - 1;
                inbox[p] = make_uint4(0u, 0u, 0u, 0u);
            }
        }
#pragma unroll
        for (int k = 0; k < K; k++) {
            const int p = k * MD2S_T + tid;
            fx[k] -= f0[k].x * f0[k].w; fy[k] -= f0[k].y * f0[k].w; fz[k] -= f0[k].z * f0[k].w;
            uint32_t m = mk[k];
            const float4* box = gf + ((size_t)p << 4);
            while (m) {
                float4 f[4];
#pragma unroll
                for (int b = 0; b < 4; b++) {
                    f[b] = make_float4(0.f, 0.f, 0.f, 0.f);
                    if (m) { f[b] = __ldcg(box + (__ffs(m) - 1)); m &= m - 1; }
                }
#pragma unroll
                for (int b = 0; b < 4; b++) { fx[k] -= f[b].x * f[b].w; fy[k] -= f[b].y * f[b].w; fz[k] -= f[b].z * f[b].w; }
            }
            if (!MOVES(k)) { fx[k] = 0.f; fy[k] = 0.f; fz[k] = 0.f; }   // frozen / empty
        }
        if (PROF) t_recv += tick() - tr0;
    };

    // ------------------------------------------------------------------------------------
    rebuild();
    const float dt = (float)pr.timestep, dtf = 0.5f * dt;
    const float dof = 3.f * (float)d->n_nonfixed - 3.f;
    float last_temp = 0.f;
    int dbg_pass = 0;

    for (int st = (P.debug_stage >= 0 ? P.debug_stage : 0); st < pr.n_stage; st++) {
        set_stage(pr.evfactor[st], pr.envfactor[st]);
        for (int sub = (pr.relax_mdsteps > 0 ? 0 : 1); sub < 2; sub++) {
            const int nsteps = sub == 0 ? pr.relax_mdsteps : pr.mdsteps[st];
            const float t0 = sub == 0 ? (float)pr.relax_temperature : (float)pr.tstart[st];
            const float t1 = sub == 0 ? (float)pr.relax_temperature : (float)pr.tstop[st];
            const float xmax = sub == 0 ? (float)pr.relax_max_velocity : (float)pr.max_velocity;
            const float vlim2 = (xmax / dt) * (xmax / dt);
            // ---- velocity nonfixed create t0 seed+st (dist uniform, mom yes, loop all) ----
            {
                uint32_t seed = (uint32_t)((d->seed + st) % 2147483647ll);
                if (seed == 0) seed = 1;
                float s0 = 0.f, s1 = 0.f, s2 = 0.f;
#pragma unroll
                for (int k = 0; k < K; k++) {
                    float a = 0.f, b = 0.f, c = 0.f;
                    if (MOVES(k)) {
                        uint32_t r = ranpark_jump(seed, 3u * (uint32_t)perm[k * MD2S_T + tid] + 1u);
                        const double AM = 1.0 / 2147483647.0;
                        a = (float)(AM * r - 0.5);
                        r = (uint32_t)(((uint64_t)r * 16807ull) % 2147483647ull); b = (float)(AM * r - 0.5);
                        r = (uint32_t)(((uint64_t)r * 16807ull) % 2147483647ull); c = (float)(AM * r - 0.5);
                    }
                    vx[k] = a; vy[k] = b; vz[k] = c;
                    s0 += a; s1 += b; s2 += c;
                }
                s0 = block_sum1(s0); s1 = block_sum1(s1); s2 = block_sum1(s2);
                const float inv = d->n_nonfixed > 0 ? 1.f / (float)d->n_nonfixed : 0.f;
                float ke = 0.f;
#pragma unroll
                for (int k = 0; k < K; k++) {
                    if (MOVES(k)) {
                        vx[k] -= s0 * inv; vy[k] -= s1 * inv; vz[k] -= s2 * inv;
                        ke += vx[k] * vx[k] + vy[k] * vy[k] + vz[k] * vz[k];
                    }
                }
                ke = block_sum1(ke);
                const float tcur = dof > 0.f ? ke / dof : 0.f;
                const float fac = tcur > 0.f ? sqrtf(t0 / tcur) : 0.f;
#pragma unroll
                for (int k = 0; k < K; k++) { vx[k] *= fac; vy[k] *= fac; vz[k] *= fac; }
            }
            // ---- run nsteps: velocity Verlet with the nve/limit speed cap + temp/rescale; step 0 = run setup ----
            float scale = 1.f;
            for (int step = 0; step <= nsteps; step++) {
                const long long tc0 = tick();
                if (bond_build) forces(std::true_type{}); else forces(std::false_type{});
                if (PROF) t_force += tick() - tc0;
                if (P.debug_stage >= 0) {       // forces of one run-setup evaluation, for the parity test of THIS force routine:
                    if (dbg_pass++ == 0) { step = -1; __syncthreads(); continue; }   // once through the full bond lists, once through the candidates
#pragma unroll
                    for (int k = 0; k < K; k++) {
                        const uint16_t a = perm[k * MD2S_T + tid];
                        if (a != MD2_NONE && V.dbg_forces) {
                            double* o = V.dbg_forces + 3 * (size_t)(d->atom_off + a);
                            o[0] = fx[k]; o[1] = fy[k]; o[2] = fz[k];
                        }
                    }
                    if (tid == 0) B.flags[blockIdx.x] = overflow ? 1 : 0;
                    return;
                }
                if (step > 0) {
                    float ke = 0.f;
#pragma unroll
                    for (int k = 0; k < K; k++) {       // second half kick (frozen/empty: f = v = 0)
                        float a = vx[k] + dtf * fx[k], b = vy[k] + dtf * fy[k], c = vz[k] + dtf * fz[k];
                        const float vsq = a * a + b * b + c * c;
                        if (vsq > vlim2) { const float sc = sqrtf(vlim2 / vsq); a *= sc; b *= sc; c *= sc; }
                        vx[k] = a; vy[k] = b; vz[k] = c;
                        ke += a * a + b * b + c * c;
                    }
                    ke = block_sum1(ke);
                    const float tcur = dof > 0.f ? ke / dof : 0.f;
                    const float ttar = t0 + ((float)step / (float)nsteps) * (t1 - t0);
                    scale = 1.f;
                    if (tcur > 0.f && fabsf(tcur - ttar) > 0.1f) scale = sqrtf(ttar / tcur);
                    last_temp = tcur * scale * scale;
                }
                if (step == nsteps) {
#pragma unroll
                    for (int k = 0; k < K; k++) { vx[k] *= scale; vy[k] *= scale; vz[k] *= scale; }
                    break;
                }
                int need = 0;
#pragma unroll
                for (int k = 0; k < K; k++) {           // rescale + first half kick + drift
                    if (MOVES(k)) {
                        const int p = k * MD2S_T + tid;
                        float a = vx[k] * scale + dtf * fx[k], b = vy[k] * scale + dtf * fy[k], c = vz[k] * scale + dtf * fz[k];
                        const float vsq = a * a + b * b + c * c;
                        if (vsq > vlim2) { const float sc = sqrtf(vlim2 / vsq); a *= sc; b *= sc; c *= sc; }
                        vx[k] = a; vy[k] = b; vz[k] = c;
                        float4 q = x[p];
                        q.x += dt * a; q.y += dt * b; q.z += dt * c;
                        x[p] = q;
                        const float ddx = q.x - xbx[p], ddy = q.y - xby[p], ddz = q.z - xbz[p];
                        need |= (ddx * ddx + ddy * ddy + ddz * ddz > lim);
                    }
                }
                if (__syncthreads_or(need)) rebuild();
            }
        }
    }
    // results: coordinates back in the original atom order (float32 like Particle.pos)
    int bad = 0;
#pragma unroll
    for (int k = 0; k < K; k++) {
        const uint16_t a = perm[k * MD2S_T + tid];
        if (a != MD2_NONE) {
            const float4 q = x[k * MD2S_T + tid];
            xg[a] = q;
            bad |= !(isfinite(q.x) && isfinite(q.y) && isfinite(q.z));
        }
    }
    bad = __syncthreads_or(bad);
    if (tid == 0) {
        igmk_result_t* r = B.res + blockIdx.x;
        const double dof_all = 3.0 * n - 3.0;     // thermo `temp` is over group all
        r->temperature = dof_all > 0 ? (double)last_temp * (double)dof / dof_all : 0.0;
        r->n_rebuilds = n_rebuild;
        B.flags[blockIdx.x] = (overflow ? 1 : 0) | (bad ? 2 : 0);
        if (PROF && V.prof) {
            long long* o = V.prof + 16 * (size_t)blockIdx.x;
            o[0] = t_force; o[1] = t_search; o[2] = t_build; o[3] = n_rebuild;
            o[4] = t_recv; o[5] = t_pair; o[6] = t_bondA; o[7] = t_smax; o[8] = t_ssum; o[9] = t_fmax; o[10] = t_fsum; o[11] = n_cand; o[12] = n_full; o[13] = n_nbr;
        }
    }
}

#undef MOVES

}  // namespace igmk
