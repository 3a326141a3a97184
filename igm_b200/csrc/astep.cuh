// astep.cuh -- A-step device code (Hi-C activation distances + contact assignment).
//
// k_actdist restates get_actdist (igm/steps/ActivationDistanceStep.py:319-459) bit for bit:
// float32 squared distances with the association ((dx^2+dy^2)+dz^2) and no FMA contraction,
// float64 probability correction (cleanProbability, :306-315), an exact order statistic in
// place of np.sort (:425,:446), float64 sqrt, and the "%10.4f"/"%.4f" text round trip + float32
// parse of :39/:241 done arithmetically.
// k_assign restates the predicate of inter_hic.py:53 / intra_hic.py:54 with the arithmetic of
// np.linalg.norm on a float32 3-vector (OpenBLAS sdot: float32 products, double accumulator).
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

namespace igmk {

constexpr int AD_THREADS = 256;
constexpr int AD_MAX_COPIES = 4;          // copies per haploid bin (diploid = 2)

struct PopView {
    int32_t n_bead, n_struct, n_hap;
    const float*   crd;        // [n_bead][n_struct][3]
    const float*   radii;      // [n_bead]
    const int32_t* chrom;      // [n_bead]
    const int32_t* copy_off;   // [n_hap+1]
    const int32_t* copy_beads;
};

// "%.{4}f" % x followed by float(str): the decimal string is the exact value of x rounded
// half-to-even to 4 decimals (glibc printf is exact), and parsing it gives the double nearest to
// r / 10^4 -- which is what IEEE division of the two exactly representable numbers returns.
// Then np.genfromtxt casts that double to float32.
__device__ __forceinline__ float text_roundtrip(double x, double scale /* 10^ndec, exact */)
{
    const double p = x * scale;                       // rounded product
    const double err = fma(x, scale, -p);             // exact residual: x*scale = p + err
    double r = rint(p);                               // half-to-even on p
    const double fl = floor(p);
    if (p - fl == 0.5) {                              // p is an exact tie: the residual decides
        if (err > 0.0) r = fl + 1.0;
        else if (err < 0.0) r = fl;
    }
    return (float)(r / scale);
}
__device__ __forceinline__ float text_roundtrip4(double x) { return text_roundtrip(x, 1.0e4); }

// order-preserving key for non-negative floats is the bit pattern itself
__device__ __forceinline__ uint32_t fkey(float v) { return __float_as_uint(v); }

// One step of the radix-256 order-statistic descent, shared by the two activation-distance kernels.
// hist[bin] = number of live keys whose current digit is `bin` (filled with shared-memory atomics, one bin
// per thread of the first AD_THREADS == 256).  Finds the bin that holds rank k, -> (digit, rank inside the bin, bin size)
// for every thread, and leaves hist zeroed for its next use.  Two barriers.
__device__ __forceinline__ void radix_pick(uint32_t* hist, uint32_t* wsum, uint32_t* pick, int tid, int lane, int warp,
                                           uint32_t& k, uint32_t& digit, int& remaining)
{
    uint32_t h = 0;
    if (tid < AD_THREADS) { h = hist[tid]; hist[tid] = 0; }    // CTAs of more than 256 threads: the others only keep the barriers
    uint32_t incl = h;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { const uint32_t t = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += t; }
    if (lane == 31 && warp < AD_THREADS / 32) wsum[warp] = incl;
    __syncthreads();
#pragma unroll
    for (int w = 0; w < AD_THREADS / 32 - 1; w++) incl += (w < warp) ? wsum[w] : 0u;
    const uint32_t excl = incl - h;
    if (k >= excl && k < incl) { pick[0] = (uint32_t)tid; pick[1] = k - excl; pick[2] = h; }   // exactly one thread: k < sum of hist, h > 0
    __syncthreads();
    digit = pick[0]; k = pick[1]; remaining = (int)pick[2];
}

// One CTA of NT threads per haploid pair.  Shared memory holds the n_comb * n_struct squared distances; when
// that leaves room for fewer than three 256-thread CTAs per SM the launcher takes NT = 1024, so that enough
// coordinate loads are in flight to stream from HBM (the slab no longer fits L2 at these sizes).
// Per-pair outputs: valid flag, activation distance and corrected probability (both already
// through the text round trip), number of copy combinations emitted.
template <int NT>
__global__ void __launch_bounds__(NT)
k_actdist(PopView P, int64_t n_pairs, const int32_t* __restrict__ pair_i, const int32_t* __restrict__ pair_j,
          const double* __restrict__ pwish, const double* __restrict__ plast, float contact_range_f32,
          float* __restrict__ out_dist, float* __restrict__ out_prob, uint8_t* __restrict__ out_valid)
{
    extern __shared__ __align__(16) unsigned char smem_raw[];
    float* dsq = (float*)smem_raw;
    __shared__ int s_red[2][NT / 32];
    __shared__ uint32_t s_mm[2][NT / 32];
    __shared__ uint32_t s_fin[32];
    __shared__ int s_nfin;
    __shared__ uint32_t s_hist[2][AD_THREADS], s_wsum[2][AD_THREADS / 32], s_pick[2][3];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int S = P.n_struct;
    const bool vec = (S & 3) == 0 && (reinterpret_cast<uintptr_t>(P.crd) & 15) == 0;     // whole groups of four, 16-byte aligned rows
    if (tid < AD_THREADS) { s_hist[0][tid] = 0; s_hist[1][tid] = 0; }   // radix_pick leaves them zeroed again
    __syncthreads();

    for (int64_t pidx = blockIdx.x; pidx < n_pairs; pidx += gridDim.x) {
        const int hi = pair_i[pidx], hj = pair_j[pidx];
        if (hi == hj) {                                   // :351-352
            if (tid == 0) out_valid[pidx] = 0;
            continue;
        }
        const int io = P.copy_off[hi], jo = P.copy_off[hj];
        const int ni = P.copy_off[hi + 1] - io, nj = P.copy_off[hj + 1] - jo;
        const bool intra = P.chrom[hi] == P.chrom[hj];
        const int ncomb = intra ? min(ni, nj) : ni * nj;  // n_possible_contacts, :384-385,:400-401
        const float ri = P.radii[P.copy_beads[io]], rj = P.radii[P.copy_beads[jo]];
        const float rc = __fmul_rn(contact_range_f32, __fadd_rn(ri, rj));
        const float rcutsq = __fmul_rn(rc, rc);           // :368, float32 under numpy >= 2
        const int total = ncomb * S;

        // ---- squared distances (A1) + contact count (A2) + min/max key ----
        int cnt = 0;
        uint32_t kmin = 0xffffffffu, kmax = 0u;
        for (int c = 0; c < ncomb; c++) {
            const int bi = P.copy_beads[io + (intra ? c : c / nj)];
            const int bj = P.copy_beads[jo + (intra ? c : c % nj)];
            const float* xi = P.crd + (size_t)bi * S * 3;
            const float* xj = P.crd + (size_t)bj * S * 3;
            if (vec) {
                // four structures at a time: 12 consecutive floats of either bead row = three 16-byte loads each
                // (a warp reads 1.5 KB contiguous per row), one 16-byte store of the four squared distances
                const float4* ri4 = reinterpret_cast<const float4*>(xi);
                const float4* rj4 = reinterpret_cast<const float4*>(xj);
                for (int g4 = tid; g4 < (S >> 2); g4 += NT) {
                    const float4 a0 = __ldg(ri4 + 3 * g4), a1 = __ldg(ri4 + 3 * g4 + 1), a2 = __ldg(ri4 + 3 * g4 + 2);
                    const float4 b0 = __ldg(rj4 + 3 * g4), b1 = __ldg(rj4 + 3 * g4 + 1), b2 = __ldg(rj4 + 3 * g4 + 2);
                    const float pa[12] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w, a2.x, a2.y, a2.z, a2.w};
                    const float pb[12] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w, b2.x, b2.y, b2.z, b2.w};
                    float v[4];
#pragma unroll
                    for (int w = 0; w < 4; w++) {
                        const float dx = __fsub_rn(pa[3 * w], pb[3 * w]);
                        const float dy = __fsub_rn(pa[3 * w + 1], pb[3 * w + 1]);
                        const float dz = __fsub_rn(pa[3 * w + 2], pb[3 * w + 2]);
                        v[w] = __fadd_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)), __fmul_rn(dz, dz));
                        cnt += (v[w] <= rcutsq);
                        const uint32_t k = fkey(v[w]);
                        kmin = min(kmin, k); kmax = max(kmax, k);
                    }
                    *reinterpret_cast<float4*>(dsq + c * S + 4 * g4) = make_float4(v[0], v[1], v[2], v[3]);
                }
            } else {
                for (int s = tid; s < S; s += NT) {
                    const float dx = __fsub_rn(xi[3 * s], xj[3 * s]);
                    const float dy = __fsub_rn(xi[3 * s + 1], xj[3 * s + 1]);
                    const float dz = __fsub_rn(xi[3 * s + 2], xj[3 * s + 2]);
                    const float v = __fadd_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)), __fmul_rn(dz, dz));
                    dsq[c * S + s] = v;
                    cnt += (v <= rcutsq);
                    const uint32_t k = fkey(v);
                    kmin = min(kmin, k); kmax = max(kmax, k);
                }
            }
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            cnt += __shfl_xor_sync(0xffffffffu, cnt, o);
            kmin = min(kmin, __shfl_xor_sync(0xffffffffu, kmin, o));
            kmax = max(kmax, __shfl_xor_sync(0xffffffffu, kmax, o));
        }
        if (lane == 0) { s_red[0][warp] = cnt; s_mm[0][warp] = kmin; s_mm[1][warp] = kmax; }
        __syncthreads();
        cnt = 0; kmin = 0xffffffffu; kmax = 0u;
#pragma unroll
        for (int w = 0; w < NT / 32; w++) { cnt += s_red[0][w]; kmin = min(kmin, s_mm[0][w]); kmax = max(kmax, s_mm[1][w]); }

        // ---- probability correction (A4), float64 exactly as the host does it ----
        const double pnow = (double)cnt / (double)(ncomb * S);                  // :422
        const double pl = plast[pidx], pw = pwish[pidx];
        double t = pl < 1.0 ? (pnow - pl) / (1.0 - pl) : pnow; t = t > 0.0 ? t : 0.0;   // :430
        double p = t < 1.0 ? (pw - t) / (1.0 - t) : pw; p = p > 0.0 ? p : 0.0;          // :434
        if (!(p > 0.0)) {
            if (tid == 0) out_valid[pidx] = 0;
            __syncthreads();
            continue;
        }
        // o = min(n*S-1, int(round(n*p*S))), product left to right, round half to even  (:443-444)
        const double prod = ((double)ncomb * p) * (double)S;
        long long o = (long long)rint(prod);
        if (o > (long long)total - 1) o = total - 1;

        // ---- exact o-th smallest (A3): radix-256 descent over (bit pattern - smallest bit pattern), most
        //      significant digit first; stops once the bin holding the wanted rank has at most 32 members,
        //      which one warp then ranks directly.  Two or three passes instead of one per bit.
        uint32_t prefix = 0, k = (uint32_t)o;
        int remaining = total;
        int bit = kmin == kmax ? -1 : 31 - __clz(kmax - kmin);        // highest bit any offset key has set
        int par = 0;
        if (tid == 0) s_nfin = 0;
        while (bit >= 0 && remaining > 32) {
            const int shift = max(bit - 7, 0);
            const uint32_t hmask = (bit == 31) ? 0u : (0xffffffffu << (bit + 1));
            for (int q = tid; q < total; q += NT) {
                const uint32_t key = fkey(dsq[q]) - kmin;
                if ((key & hmask) == prefix) atomicAdd(&s_hist[par][(key - prefix) >> shift], 1u);
            }
            __syncthreads();
            uint32_t digit;
            radix_pick(s_hist[par], s_wsum[par], s_pick[par], tid, lane, warp, k, digit, remaining);
            prefix |= digit << shift;
            bit = shift - 1;
            par ^= 1;
        }
        if (bit >= 0) {
            const uint32_t hmask = (bit == 31) ? 0u : (0xffffffffu << (bit + 1));
            __syncthreads();                              // s_nfin = 0 is visible
            for (int q = tid; q < total; q += NT) {
                const uint32_t key = fkey(dsq[q]) - kmin;
                if ((key & hmask) == prefix) s_fin[atomicAdd(&s_nfin, 1)] = key;
            }
            __syncthreads();
            if (warp == 0) {
                const uint32_t mine = lane < remaining ? s_fin[lane] : 0xffffffffu;
                int rank = 0;
#pragma unroll
                for (int m = 0; m < 32; m++) {
                    const uint32_t other = __shfl_sync(0xffffffffu, mine, m);
                    rank += (other < mine) || (other == mine && m < lane);
                }
                const uint32_t hit = __ballot_sync(0xffffffffu, rank == (int)k && lane < remaining);
                prefix = __shfl_sync(0xffffffffu, mine, __ffs(hit) - 1);
            }
        }
        prefix += kmin;                                   // back to the bit pattern of the squared distance
        if (tid == 0) {
            const double ad = sqrt((double)__uint_as_float(prefix));            // :446
            out_dist[pidx] = text_roundtrip4(ad);
            out_prob[pidx] = text_roundtrip4(p);
            out_valid[pidx] = 1;
        }
        __syncthreads();
    }
}

// GW consecutive structures of one bead row (3 GW floats at `row`), of which the first `nvalid` exist
template <int GW>
__device__ __forceinline__ void load_group(const float* __restrict__ row, int nvalid, bool vec, float (&v)[3 * GW])
{
    if constexpr (GW == 4) {
        if (vec) {                                        // whole groups only: n_struct is a multiple of 4
            if (nvalid > 0) {
                const float4* r4 = reinterpret_cast<const float4*>(row);
                const float4 q0 = __ldg(r4), q1 = __ldg(r4 + 1), q2 = __ldg(r4 + 2);
                v[0] = q0.x; v[1] = q0.y; v[2] = q0.z; v[3] = q0.w; v[4] = q1.x; v[5] = q1.y; v[6] = q1.z; v[7] = q1.w;
                v[8] = q2.x; v[9] = q2.y; v[10] = q2.z; v[11] = q2.w;
            }
            return;
        }
    }
#pragma unroll
    for (int w = 0; w < GW; w++)
        if (w < nvalid) { v[3 * w] = __ldg(row + 3 * w); v[3 * w + 1] = __ldg(row + 3 * w + 1); v[3 * w + 2] = __ldg(row + 3 * w + 2); }
}

// Register-resident variant for populations of up to 256 * SPT structures (n_comb <= 4): every thread
// keeps its 4 * SPT squared distances in registers, the radix-256 descent stops as soon as the bin
// that holds the wanted order statistic has at most 32 members, and one warp ranks those directly.
// Same arithmetic, same results as k_actdist; roughly 4x fewer instructions per pair.
// Everything k_actdist_reg needs to know about one pair, resolved once by k_actdist_meta: the chain
// pair -> copy_off -> copy_beads -> radii is four dependent loads, too long to sit in front of every pair.
struct __align__(16) PairMeta {
    int32_t bi[4], bj[4];        // beads of combination c (copies of i and of j)
    float   rcutsq;              // (contact_range * (ri + rj))^2 in float32, :368
    int32_t ncomb;               // n_possible_contacts; 0: i == j, no row (:351-352)
    int32_t pad[2];
    double  pl, pw;              // plast, pwish
};
static_assert(sizeof(PairMeta) == 64, "PairMeta is four 16-byte loads");

__global__ void __launch_bounds__(256)
k_actdist_meta(PopView P, int64_t n_pairs, const int32_t* __restrict__ pair_i, const int32_t* __restrict__ pair_j,
               const double* __restrict__ pwish, const double* __restrict__ plast, float contact_range_f32,
               PairMeta* __restrict__ meta)
{
    for (int64_t pidx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; pidx < n_pairs; pidx += (int64_t)gridDim.x * blockDim.x) {
        const int hi = pair_i[pidx], hj = pair_j[pidx];
        PairMeta m;
        for (int c = 0; c < 4; c++) { m.bi[c] = 0; m.bj[c] = 0; }
        m.rcutsq = 0.f; m.ncomb = 0; m.pad[0] = m.pad[1] = 0; m.pl = plast[pidx]; m.pw = pwish[pidx];
        if (hi != hj) {
            const int io = P.copy_off[hi], jo = P.copy_off[hj];
            const int ni = P.copy_off[hi + 1] - io, nj = P.copy_off[hj + 1] - jo;
            const bool intra = P.chrom[hi] == P.chrom[hj];
            m.ncomb = min(intra ? min(ni, nj) : ni * nj, 4);      // :384-385, :400-401 (the launcher guarantees <= 4)
            for (int c = 0; c < m.ncomb; c++) {
                m.bi[c] = P.copy_beads[io + (intra ? c : c / nj)];
                m.bj[c] = P.copy_beads[jo + (intra ? c : c % nj)];
            }
            const float ri = P.radii[P.copy_beads[io]], rj = P.radii[P.copy_beads[jo]];
            const float rc = __fmul_rn(contact_range_f32, __fadd_rn(ri, rj));
            m.rcutsq = __fmul_rn(rc, rc);
        }
        meta[pidx] = m;
    }
}

template <int SPT, int MINB>
__global__ void __launch_bounds__(AD_THREADS, MINB)
k_actdist_reg(PopView P, int64_t n_pairs, const PairMeta* __restrict__ meta,
              float* __restrict__ out_dist, float* __restrict__ out_prob, uint8_t* __restrict__ out_valid)
{
    // buffers indexed by the parity of the pair: no barrier at the end of a pair, the warps that do not
    // take part in the final ranking go on to the loads of the next pair
    __shared__ int s_red[2][AD_THREADS / 32];
    __shared__ uint32_t s_mm[2][2][AD_THREADS / 32];
    __shared__ uint32_t s_fin[2][32];
    __shared__ int s_nfin[2];
    __shared__ uint32_t s_hist[2][AD_THREADS], s_wsum[2][AD_THREADS / 32], s_pick[2][3];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int S = P.n_struct;
    s_hist[0][tid] = 0; s_hist[1][tid] = 0;               // radix_pick leaves them zeroed again
    __syncthreads();
    // Thread t owns the structures GW * (t + 256 g) .. + GW - 1 of group g: GW consecutive structures are
    // 3 GW consecutive floats of a bead row, fetched with three 16-byte loads when GW == 4 and the rows are
    // 16-byte aligned (a warp still reads one contiguous run).
    constexpr int GW = SPT < 4 ? SPT : 4, NG = SPT / GW;
    const bool vec = GW == 4 && (S & 3) == 0 && (reinterpret_cast<uintptr_t>(P.crd) & 15) == 0;
    int pp = 0;

    for (int64_t pidx = blockIdx.x; pidx < n_pairs; pidx += gridDim.x) {
        // the record is read piecewise where it is needed (L1 hits after the first touch): fewer live registers
        const PairMeta* mp = meta + pidx;
        if (pidx + gridDim.x < n_pairs && tid == 0)
            asm volatile("prefetch.global.L2 [%0];" ::"l"(mp + gridDim.x));
        const int2 head = __ldg(reinterpret_cast<const int2*>(&mp->rcutsq));
        const int ncomb = head.y;
        if (ncomb == 0) {                                 // :351-352
            if (tid == 0) out_valid[pidx] = 0;
            continue;
        }
        const float rcutsq = __int_as_float(head.x);
        const int total = ncomb * S;

        uint32_t key[4 * SPT];                            // 0xffffffff: empty slot (never matches a prefix)
#pragma unroll
        for (int u = 0; u < 4 * SPT; u++) key[u] = 0xffffffffu;
        int cnt = 0;
        uint32_t kmin = 0xffffffffu, kmax = 0u;
#pragma unroll
        for (int g = 0; g < NG; g++) {
            const int sbase = GW * (tid + AD_THREADS * g);
            if (sbase < S) {
                float xi[3 * GW], xj[3 * GW];
                int bi_prev = -1;
#pragma unroll
                for (int c = 0; c < 4; c++) {
                    if (c < ncomb) {
                        const int bic = __ldg(&mp->bi[c]), bjc = __ldg(&mp->bj[c]);
                        if (c == 0 || bic != bi_prev)     // a copy of i that serves two combinations is loaded once
                            load_group<GW>(P.crd + ((size_t)bic * S + sbase) * 3, S - sbase, vec, xi);
                        bi_prev = bic;
                        load_group<GW>(P.crd + ((size_t)bjc * S + sbase) * 3, S - sbase, vec, xj);
#pragma unroll
                        for (int w = 0; w < GW; w++) {
                            if (vec || sbase + w < S) {   // vec: whole groups only
                                const float dx = __fsub_rn(xi[3 * w], xj[3 * w]);
                                const float dy = __fsub_rn(xi[3 * w + 1], xj[3 * w + 1]);
                                const float dz = __fsub_rn(xi[3 * w + 2], xj[3 * w + 2]);
                                const float v = __fadd_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)), __fmul_rn(dz, dz));
                                cnt += (v <= rcutsq);
                                const uint32_t k = fkey(v);
                                kmin = min(kmin, k); kmax = max(kmax, k);
                                key[c * SPT + g * GW + w] = k;
                            }
                        }
                    }
                }
            }
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            cnt += __shfl_xor_sync(0xffffffffu, cnt, o);
            kmin = min(kmin, __shfl_xor_sync(0xffffffffu, kmin, o));
            kmax = max(kmax, __shfl_xor_sync(0xffffffffu, kmax, o));
        }
        if (lane == 0) { s_red[pp][warp] = cnt; s_mm[pp][0][warp] = kmin; s_mm[pp][1][warp] = kmax; }
        if (tid == 0) s_nfin[pp] = 0;
        __syncthreads();
        cnt = 0; kmin = 0xffffffffu; kmax = 0u;
#pragma unroll
        for (int w = 0; w < AD_THREADS / 32; w++) { cnt += s_red[pp][w]; kmin = min(kmin, s_mm[pp][0][w]); kmax = max(kmax, s_mm[pp][1][w]); }
        const int pq = pp;
        pp ^= 1;

        const double pnow = (double)cnt / (double)(ncomb * S);                  // :422
        const double pl = __ldg(&mp->pl), pw = __ldg(&mp->pw);
        double t = pl < 1.0 ? (pnow - pl) / (1.0 - pl) : pnow; t = t > 0.0 ? t : 0.0;   // :430
        double p = t < 1.0 ? (pw - t) / (1.0 - t) : pw; p = p > 0.0 ? p : 0.0;          // :434
        if (!(p > 0.0)) {
            if (tid == 0) out_valid[pidx] = 0;
            continue;
        }
        const double prod = ((double)ncomb * p) * (double)S;                    // :443-444
        long long o = (long long)rint(prod);
        if (o > (long long)total - 1) o = total - 1;

        // exact o-th smallest: radix-256 descent over (bit pattern - smallest bit pattern), as in k_actdist
#pragma unroll
        for (int u = 0; u < 4 * SPT; u++) key[u] -= kmin;         // empty slots keep bit 31: kmin < 2^31
        uint32_t prefix = 0, k = (uint32_t)o;
        int remaining = total;
        int bit = kmin == kmax ? -1 : 31 - __clz(kmax - kmin);        // <= 30: the keys are bit patterns of floats >= 0
        int par = 0;
        while (bit >= 0 && remaining > 32) {
            const int shift = max(bit - 7, 0);
            const uint32_t hmask = 0xffffffffu << (bit + 1);
#pragma unroll
            for (int u = 0; u < 4 * SPT; u++)
                if ((key[u] & hmask) == prefix) atomicAdd(&s_hist[par][(key[u] - prefix) >> shift], 1u);
            __syncthreads();
            uint32_t digit;
            radix_pick(s_hist[par], s_wsum[par], s_pick[par], tid, lane, warp, k, digit, remaining);
            prefix |= digit << shift;
            bit = shift - 1;
            par ^= 1;
        }
        uint32_t answer = prefix;
        if (bit >= 0) {
            // <= 32 keys share `prefix` above `bit`: collect them, rank them in one warp
            const uint32_t hmask = 0xffffffffu << (bit + 1);
#pragma unroll
            for (int u = 0; u < 4 * SPT; u++)
                if ((key[u] & hmask) == prefix) s_fin[pq][atomicAdd(&s_nfin[pq], 1)] = key[u];
            __syncthreads();
            if (warp == 0) {
                const uint32_t mine = lane < remaining ? s_fin[pq][lane] : 0xffffffffu;
                int rank = 0;
#pragma unroll
                for (int mm = 0; mm < 32; mm++) {
                    const uint32_t other = __shfl_sync(0xffffffffu, mine, mm);
                    rank += (other < mine) || (other == mine && mm < lane);
                }
                const uint32_t hit = __ballot_sync(0xffffffffu, rank == (int)k && lane < remaining);
                answer = __shfl_sync(0xffffffffu, mine, __ffs(hit) - 1);
            }
        }
        if (tid == 0) {
            const double ad = sqrt((double)__uint_as_float(answer + kmin));     // :446; + kmin: back to the bit pattern
            out_dist[pidx] = text_roundtrip4(ad);
            out_prob[pidx] = text_roundtrip4(p);
            out_valid[pidx] = 1;
        }
    }
}

// Contact frequency of haploid pairs in the population: the `pnow` of get_actdist
// (ActivationDistanceStep.py:419-422) = count(d_sq <= rcutsq) / (n_comb * n_struct), as float32.
// One warp per pair.  Used to derive geometrically consistent synthetic Hi-C matrices and as the
// population-side half of a Hi-C evaluation.
__global__ void __launch_bounds__(256)
k_contact_freq(PopView P, int64_t n_pairs, const int32_t* __restrict__ pair_i, const int32_t* __restrict__ pair_j,
               float contact_range_f32, float* __restrict__ freq)
{
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int S = P.n_struct;
    for (int64_t pidx = (int64_t)blockIdx.x * 8 + warp; pidx < n_pairs; pidx += (int64_t)gridDim.x * 8) {
        const int hi = pair_i[pidx], hj = pair_j[pidx];
        const int io = P.copy_off[hi], jo = P.copy_off[hj];
        const int ni = P.copy_off[hi + 1] - io, nj = P.copy_off[hj + 1] - jo;
        const bool intra = P.chrom[hi] == P.chrom[hj];
        const int ncomb = intra ? min(ni, nj) : ni * nj;
        const float rc = __fmul_rn(contact_range_f32, __fadd_rn(P.radii[P.copy_beads[io]], P.radii[P.copy_beads[jo]]));
        const float rcutsq = __fmul_rn(rc, rc);
        int cnt = 0;
        for (int c = 0; c < ncomb; c++) {
            const float* xi = P.crd + (size_t)P.copy_beads[io + (intra ? c : c / nj)] * S * 3;
            const float* xj = P.crd + (size_t)P.copy_beads[jo + (intra ? c : c % nj)] * S * 3;
            for (int s = lane; s < S; s += 32) {
                const float dx = __fsub_rn(xi[3 * s], xj[3 * s]);
                const float dy = __fsub_rn(xi[3 * s + 1], xj[3 * s + 1]);
                const float dz = __fsub_rn(xi[3 * s + 2], xj[3 * s + 2]);
                cnt += (__fadd_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)), __fmul_rn(dz, dz)) <= rcutsq);
            }
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) cnt += __shfl_xor_sync(0xffffffffu, cnt, o);
        if (lane == 0) freq[pidx] = hi == hj ? 1.0f : (float)((double)cnt / (double)(ncomb * S));
    }
}

// Contact assignment.  One warp per actdist row, lanes over structures; 32 rows per CTA.
// bits[r][w] bit l  <=>  structure 32*w+l gets a bond for row r.
__global__ void __launch_bounds__(1024)
k_assign(PopView P, int64_t n_rows, const int32_t* __restrict__ row, const int32_t* __restrict__ col,
         const float* __restrict__ dist, uint32_t* __restrict__ bits)
{
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int S = P.n_struct, nw = (S + 31) >> 5;
    for (int64_t r = (int64_t)blockIdx.x * 32 + warp; r < n_rows; r += (int64_t)gridDim.x * 32) {
        const float* xi = P.crd + (size_t)row[r] * S * 3;
        const float* xj = P.crd + (size_t)col[r] * S * 3;
        const float d = dist[r];
        for (int w = 0; w < nw; w++) {
            const int s = w * 32 + lane;
            bool hit = false;
            if (s < S) {
                const float dx = __fsub_rn(xi[3 * s], xj[3 * s]);
                const float dy = __fsub_rn(xi[3 * s + 1], xj[3 * s + 1]);
                const float dz = __fsub_rn(xi[3 * s + 2], xj[3 * s + 2]);
                const double acc = __dadd_rn(__dadd_rn((double)__fmul_rn(dx, dx), (double)__fmul_rn(dy, dy)),
                                             (double)__fmul_rn(dz, dz));
                hit = __fsqrt_rn(__double2float_rn(acc)) <= d;
            }
            const uint32_t b = __ballot_sync(0xffffffffu, hit);
            if (lane == 0) bits[r * nw + w] = b;
        }
    }
}

// ---------------------------------------------------------------------------------------
// DamID (lamina) A-step.  k_damid_actdist restates get_damid_actdist
// (igm/steps/DamidActivationDistanceStep.py:285-356): per locus the normalised squared distance to
// the shrunken envelope of every copy in every structure -- float32 squares, then float64
// arithmetic exactly as numpy promotes it (snormsq_sphere :35-54: float32 sum (x2+y2)+z2 divided by
// the float64 (R-r)^2; snormsq_ellipse :58-76: each float32 square divided by a float64 (axis-r)^2,
// summed left to right) -- the count of values >= 1, the two cleanProbability corrections, the
// o-th LARGEST value (descending sort, :335) and the "%10.2f" / "%.5f" text round trip (:31).
// One CTA per locus; the values live in shared memory as float64.
// The denominators (axis*(1-contact_range) - r)**2 come from the host, which evaluates them with
// libm pow(x, 2.0) like numpy's scalar `**` does (it differs from x*x in ~0.1 % of the cases).

__global__ void __launch_bounds__(AD_THREADS)
k_damid_actdist(PopView P, int64_t n_loci, const int32_t* __restrict__ loc, const double* __restrict__ pwish,
                const double* __restrict__ plast, const double* __restrict__ den /* [n_loci][3] */, int ellipsoid,
                float* __restrict__ out_dist, float* __restrict__ out_prob)
{
    extern __shared__ __align__(16) unsigned char smem_raw[];
    unsigned long long* key = (unsigned long long*)smem_raw;      // non-negative doubles order like their bit patterns
    __shared__ int s_red[2][AD_THREADS / 32];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int S = P.n_struct;
    for (int64_t q = blockIdx.x; q < n_loci; q += gridDim.x) {
        const int h = loc[q];
        const int o0 = P.copy_off[h], nc = P.copy_off[h + 1] - o0;
        const int total = nc * S;
        const double a2 = den[3 * q], b2 = den[3 * q + 1], c2 = den[3 * q + 2];
        int cnt = 0;
        for (int c = 0; c < nc; c++) {
            const float* x = P.crd + (size_t)P.copy_beads[o0 + c] * S * 3;
            for (int s = tid; s < S; s += AD_THREADS) {
                const float xx = __fmul_rn(x[3 * s], x[3 * s]), yy = __fmul_rn(x[3 * s + 1], x[3 * s + 1]),
                            zz = __fmul_rn(x[3 * s + 2], x[3 * s + 2]);
                double v;
                if (ellipsoid) v = __dadd_rn(__dadd_rn(__ddiv_rn((double)xx, a2), __ddiv_rn((double)yy, b2)), __ddiv_rn((double)zz, c2));
                else v = __ddiv_rn((double)__fadd_rn(__fadd_rn(xx, yy), zz), a2);
                key[c * S + s] = (unsigned long long)__double_as_longlong(v);
                cnt += (v >= 1.0);
            }
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) cnt += __shfl_xor_sync(0xffffffffu, cnt, o);
        if (lane == 0) s_red[0][warp] = cnt;
        __syncthreads();
        cnt = 0;
#pragma unroll
        for (int w = 0; w < AD_THREADS / 32; w++) cnt += s_red[0][w];
        const double pnow = (double)cnt / (double)S / (double)nc;                     // :339
        const double pl = plast[q], pw = pwish[q];
        double t = pl < 1.0 ? (pnow - pl) / (1.0 - pl) : pnow; t = t > 0.0 ? t : 0.0;
        double p = t < 1.0 ? (pw - t) / (1.0 - t) : pw; p = p > 0.0 ? p : 0.0;
        double ad = 2.0;                                                              // :347
        if (p > 0.0) {
            long long o = (long long)rint((double)total * p);                         // :350-351
            if (o > (long long)total - 1) o = total - 1;
            // o-th largest == (total-1-o)-th smallest; bitwise descent over the 64-bit keys
            unsigned long long prefix = 0ull;
            unsigned int k = (unsigned int)(total - 1 - o);
            int par = 1;
            for (int bit = 62; bit >= 0; bit--) {                 // bit 63 (sign) is clear in every key
                const unsigned long long hmask = ~0ull << (bit + 1);
                int c0 = 0;
                for (int i = tid; i < total; i += AD_THREADS) {
                    const unsigned long long kk = key[i];
                    c0 += ((kk & hmask) == prefix) && !((kk >> bit) & 1ull);
                }
#pragma unroll
                for (int o2 = 16; o2 > 0; o2 >>= 1) c0 += __shfl_xor_sync(0xffffffffu, c0, o2);
                if (lane == 0) s_red[par][warp] = c0;
                __syncthreads();
                c0 = 0;
#pragma unroll
                for (int w = 0; w < AD_THREADS / 32; w++) c0 += s_red[par][w];
                par ^= 1;
                if (k >= (unsigned int)c0) { k -= (unsigned int)c0; prefix |= (1ull << bit); }
            }
            ad = sqrt(__longlong_as_double((long long)prefix));                       // :354
        }
        if (tid == 0) {
            out_dist[q] = text_roundtrip(ad, 1.0e2);
            out_prob[q] = text_roundtrip(p, 1.0e5);
        }
        __syncthreads();
    }
}

// Damid._apply_envelope (igm/restraints/damid.py:116-130): bead `loc[r]` of structure s joins the
// lamina envelope iff snormsq_ellipsoid(pos, abc*cutoff, r) >= d^2, with d^2 squared in float32 and
// widened, the left side in float64 from float32 squares.  One warp per row; bit matrix like k_assign.
__global__ void __launch_bounds__(1024)
k_damid_assign(PopView P, int64_t n_rows, const int32_t* __restrict__ bead, const float* __restrict__ dist,
               const double* __restrict__ den /* [n_rows][3] */, uint32_t* __restrict__ bits)
{
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int S = P.n_struct, nw = (S + 31) >> 5;
    for (int64_t r = (int64_t)blockIdx.x * 32 + warp; r < n_rows; r += (int64_t)gridDim.x * 32) {
        const float* x = P.crd + (size_t)bead[r] * S * 3;
        const double a2 = den[3 * r], b2 = den[3 * r + 1], c2 = den[3 * r + 2];
        const double d2 = (double)__fmul_rn(dist[r], dist[r]);
        for (int w = 0; w < nw; w++) {
            const int s = w * 32 + lane;
            bool hit = false;
            if (s < S) {
                const float xx = __fmul_rn(x[3 * s], x[3 * s]), yy = __fmul_rn(x[3 * s + 1], x[3 * s + 1]),
                            zz = __fmul_rn(x[3 * s + 2], x[3 * s + 2]);
                const double v = __dadd_rn(__dadd_rn(__ddiv_rn((double)xx, a2), __ddiv_rn((double)yy, b2)), __ddiv_rn((double)zz, c2));
                hit = v >= d2;
            }
            const uint32_t b = __ballot_sync(0xffffffffu, hit);
            if (lane == 0) bits[r * nw + w] = b;
        }
    }
}

// ---------------------------------------------------------------------------------------
// FISH A-step (igm/steps/FishAssignmentStep.py): per probe pair the four diploid distances
// (get_pair_dists :31-51) or per probe the two radial distances (get_rad_dists :55-72) in every
// structure -- np.linalg.norm(axis=1) on float32 = sqrtf((x2+y2)+z2) -- their minimum and maximum
// over the copies, and the rank of every structure in the population for both
// (get_min_max_and_idx :74-88, argsort(argsort(.))).  Ties (practically absent for float32
// distances) are ranked by structure index, numpy's quicksort leaves their order unspecified.
// One CTA per item; hap_j < 0 marks a radial probe.
__global__ void __launch_bounds__(AD_THREADS)
k_fish_ranks(PopView P, int64_t n_items, const int32_t* __restrict__ hap_i, const int32_t* __restrict__ hap_j,
             float* __restrict__ out_min, float* __restrict__ out_max, int32_t* __restrict__ rank_min, int32_t* __restrict__ rank_max)
{
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int S = P.n_struct;
    float* vmin = (float*)smem_raw;
    float* vmax = vmin + S;
    for (int64_t q = blockIdx.x; q < n_items; q += gridDim.x) {
        const int hi = hap_i[q], hj = hap_j[q];
        const float* x1 = P.crd + (size_t)P.copy_beads[P.copy_off[hi]] * S * 3;
        const float* x2 = P.crd + (size_t)P.copy_beads[P.copy_off[hi] + 1] * S * 3;
        const float *y1 = nullptr, *y2 = nullptr;
        if (hj >= 0) {
            y1 = P.crd + (size_t)P.copy_beads[P.copy_off[hj]] * S * 3;
            y2 = P.crd + (size_t)P.copy_beads[P.copy_off[hj] + 1] * S * 3;
        }
        auto norm3 = [](float a, float b, float c) {
            return __fsqrt_rn(__fadd_rn(__fadd_rn(__fmul_rn(a, a), __fmul_rn(b, b)), __fmul_rn(c, c)));
        };
        for (int s = threadIdx.x; s < S; s += AD_THREADS) {
            const float ax = x1[3 * s], ay = x1[3 * s + 1], az = x1[3 * s + 2];
            const float bx = x2[3 * s], by = x2[3 * s + 1], bz = x2[3 * s + 2];
            float lo, hi2;
            if (hj >= 0) {
                const float cx = y1[3 * s], cy = y1[3 * s + 1], cz = y1[3 * s + 2];
                const float dx = y2[3 * s], dy = y2[3 * s + 1], dz = y2[3 * s + 2];
                const float d0 = norm3(__fsub_rn(ax, cx), __fsub_rn(ay, cy), __fsub_rn(az, cz));
                const float d1 = norm3(__fsub_rn(ax, dx), __fsub_rn(ay, dy), __fsub_rn(az, dz));
                const float d2 = norm3(__fsub_rn(bx, cx), __fsub_rn(by, cy), __fsub_rn(bz, cz));
                const float d3 = norm3(__fsub_rn(bx, dx), __fsub_rn(by, dy), __fsub_rn(bz, dz));
                lo = fminf(fminf(d0, d1), fminf(d2, d3)); hi2 = fmaxf(fmaxf(d0, d1), fmaxf(d2, d3));
            } else {
                const float d0 = norm3(ax, ay, az), d1 = norm3(bx, by, bz);
                lo = fminf(d0, d1); hi2 = fmaxf(d0, d1);
            }
            vmin[s] = lo; vmax[s] = hi2;
            out_min[q * S + s] = lo; out_max[q * S + s] = hi2;
        }
        __syncthreads();
        for (int s = threadIdx.x; s < S; s += AD_THREADS) {
            const float a = vmin[s], b = vmax[s];
            int ra = 0, rb = 0;
            for (int t = 0; t < S; t++) {
                const float u = vmin[t], w = vmax[t];
                ra += (u < a) || (u == a && t < s);
                rb += (w < b) || (w == b && t < s);
            }
            rank_min[q * S + s] = ra; rank_max[q * S + s] = rb;
        }
        __syncthreads();
    }
}

// ---------------------------------------------------------------------------------------
// SPRITE A-step numerics: get_rg2s_cpp (igm/cython_compiled/cpp_sprite_assignment.cpp:79-143).
// A cluster is a list of regions, each with K(i) alternative beads (chromosome copies); for every
// structure the combination of alternatives with the smallest squared radius of gyration is found
// by enumeration in the reference's order (get_combination :63-77: region 0 is the fastest digit)
// with its float32 arithmetic (gyration_radius_sq :49-61: sequential mean, divide by float(n),
// sum of ((dx*dx + dy*dy) + dz*dz), divide by float(n); strict `<` keeps the first minimum).
// One thread per (cluster, structure); a warp-aligned group of structures per cluster.
constexpr int SPRITE_MAX_REGIONS = 8;

__global__ void __launch_bounds__(256)
k_sprite_rg2(PopView P, int32_t n_clusters, const int32_t* __restrict__ cl_off /* [n_clusters+1] regions */,
             const int32_t* __restrict__ rg_off /* [n_regions+1] beads */, const int32_t* __restrict__ beads,
             float* __restrict__ rg2 /* [n_clusters][S] */, int8_t* __restrict__ copy_idx /* [n_regions][S] */,
             int32_t* __restrict__ min_struct /* [n_clusters] */)
{
    __shared__ float s_val[8];
    __shared__ int s_idx[8];
    const int S = P.n_struct;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int c = blockIdx.x; c < n_clusters; c += gridDim.x) {
        const int r0 = cl_off[c], nr = cl_off[c + 1] - r0;
        int ncopy[SPRITE_MAX_REGIONS], boff[SPRITE_MAX_REGIONS];
        int ncomb = 1;
#pragma unroll
        for (int i = 0; i < SPRITE_MAX_REGIONS; i++) {
            ncopy[i] = 1; boff[i] = 0;
            if (i < nr) { boff[i] = rg_off[r0 + i]; ncopy[i] = rg_off[r0 + i + 1] - boff[i]; ncomb *= ncopy[i]; }
        }
        const float fn = (float)nr;
        float best_v = 3.0e38f; int best_s = 0x7fffffff;
        for (int s = threadIdx.x; s < S; s += blockDim.x) {
            float min_rg2 = 100000000.0f;             // INF of the reference (:4)
            int min_k = -1;
            for (int k = 0; k < ncomb; k++) {
                float px[SPRITE_MAX_REGIONS], py[SPRITE_MAX_REGIONS], pz[SPRITE_MAX_REGIONS];
                float mx = 0.f, my = 0.f, mz = 0.f;
                int kk = k;
#pragma unroll
                for (int i = 0; i < SPRITE_MAX_REGIONS; i++) {
                    if (i < nr) {
                        const int si = kk % ncopy[i]; kk /= ncopy[i];
                        const float* x = P.crd + ((size_t)beads[boff[i] + si] * S + s) * 3;
                        px[i] = x[0]; py[i] = x[1]; pz[i] = x[2];
                        mx = __fadd_rn(mx, px[i]); my = __fadd_rn(my, py[i]); mz = __fadd_rn(mz, pz[i]);
                    }
                }
                mx = __fdiv_rn(mx, fn); my = __fdiv_rn(my, fn); mz = __fdiv_rn(mz, fn);
                float rg = 0.f;
#pragma unroll
                for (int i = 0; i < SPRITE_MAX_REGIONS; i++) {
                    if (i < nr) {
                        const float dx = __fsub_rn(px[i], mx), dy = __fsub_rn(py[i], my), dz = __fsub_rn(pz[i], mz);
                        rg = __fadd_rn(rg, __fadd_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)), __fmul_rn(dz, dz)));
                    }
                }
                rg = __fdiv_rn(rg, fn);
                if (rg < min_rg2) { min_rg2 = rg; min_k = k; }
            }
            rg2[(size_t)c * S + s] = min_rg2;
            int kk = min_k;
#pragma unroll
            for (int i = 0; i < SPRITE_MAX_REGIONS; i++) {
                if (i < nr) {
                    int si = -1;                       // min_copy_idx starts at -1 (:118) and stays there if no rg2 < INF
                    if (min_k >= 0) { si = kk % ncopy[i]; kk /= ncopy[i]; }
                    copy_idx[(size_t)(r0 + i) * S + s] = (int8_t)si;
                }
            }
            if (min_rg2 < best_v || (min_rg2 == best_v && s < best_s)) { best_v = min_rg2; best_s = s; }
        }
        // first structure with the smallest rg2 (:136-140, strict `<`, initial minimum INF)
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const float ov = __shfl_xor_sync(0xffffffffu, best_v, o);
            const int os = __shfl_xor_sync(0xffffffffu, best_s, o);
            if (ov < best_v || (ov == best_v && os < best_s)) { best_v = ov; best_s = os; }
        }
        if (lane == 0) { s_val[warp] = best_v; s_idx[warp] = best_s; }
        __syncthreads();
        if (threadIdx.x == 0) {
            for (int w = 1; w < (int)(blockDim.x >> 5); w++)
                if (s_val[w] < best_v || (s_val[w] == best_v && s_idx[w] < best_s)) { best_v = s_val[w]; best_s = s_idx[w]; }
            min_struct[c] = best_v < 100000000.0f ? best_s : -1;      // the reference leaves it unset in that case
        }
        __syncthreads();
    }
}

// SPRITE A-step, the whole of compute_gyration_radius (igm/cython_compiled/sprite.pyx:104-283) and the keep-best
// selection of SpriteAssignmentStep.task (igm/steps/SpriteAssignmentStep.py:148-160) for one cluster per CTA.
// A cluster is a list of segments (haploid ids, ascending = grouped by chromosome); segment i has the copies
// beads[seg_off[i] .. seg_off[i+1]) and belongs to group seg_group[i] (its chromosome, numbered 0.. inside the cluster).
//   * one group (sprite.pyx:198-212): every copy k of the chromosome is tried for ALL segments, the copy with the
//     smallest Rg^2 wins (first minimum);
//   * several groups (:215-283): the copy of every chromosome is chosen on ONE representative segment per group
//     (rep[g], chosen by the caller: the reference draws it with np.random.choice) by enumerating the combinations
//     like get_rg2s_cpp (region 0 = fastest digit, strict <), then Rg^2 is taken over all segments with those copies.
// float32 arithmetic in the reference's order (gyration_radius_sq, cpp_sprite_assignment.cpp:49-61).
// Then the `keep` structures with the smallest Rg^2 in ascending order (ties by structure index), their Rg^2 and,
// per segment, the bead chosen in each of them.
constexpr int SPRITE_MAX_GROUPS = 8;
constexpr int SPRITE_SORT_MAX = 4096;      // structures the shared-memory sort holds

__device__ __forceinline__ float sprite_rg2_of(const PopView& P, int s, int seg0, int nseg, const int32_t* __restrict__ seg_off,
                                               const int32_t* __restrict__ beads, const int32_t* __restrict__ seg_group, uint32_t choice)
{
    const int S = P.n_struct;
    float mx = 0.f, my = 0.f, mz = 0.f;
    for (int i = 0; i < nseg; i++) {
        const int c = (choice >> (4 * seg_group[seg0 + i])) & 15;
        const float* x = P.crd + ((size_t)beads[seg_off[seg0 + i] + c] * S + s) * 3;
        mx = __fadd_rn(mx, x[0]); my = __fadd_rn(my, x[1]); mz = __fadd_rn(mz, x[2]);
    }
    const float fn = (float)nseg;
    mx = __fdiv_rn(mx, fn); my = __fdiv_rn(my, fn); mz = __fdiv_rn(mz, fn);
    float rg = 0.f;
    for (int i = 0; i < nseg; i++) {
        const int c = (choice >> (4 * seg_group[seg0 + i])) & 15;
        const float* x = P.crd + ((size_t)beads[seg_off[seg0 + i] + c] * S + s) * 3;
        const float dx = __fsub_rn(x[0], mx), dy = __fsub_rn(x[1], my), dz = __fsub_rn(x[2], mz);
        rg = __fadd_rn(rg, __fadd_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)), __fmul_rn(dz, dz)));
    }
    return __fdiv_rn(rg, fn);
}

__global__ void __launch_bounds__(256)
k_sprite_select(PopView P, int32_t n_clusters, const int32_t* __restrict__ cl_off, const int32_t* __restrict__ seg_off,
                const int32_t* __restrict__ beads, const int32_t* __restrict__ seg_group, const int32_t* __restrict__ n_groups,
                const int32_t* __restrict__ rep /* [n_clusters][SPRITE_MAX_GROUPS] segment (within the cluster) of each group */,
                int32_t keep, float* __restrict__ best_rg2 /* [n_clusters][keep] */, int32_t* __restrict__ best_struct,
                int32_t* __restrict__ best_beads /* [n_segments][keep] */)
{
    extern __shared__ __align__(16) unsigned char sp_smem[];
    const int S = P.n_struct;
    int npow = 1;
    while (npow < S) npow <<= 1;
    float* val = (float*)sp_smem;                     // [npow]
    int32_t* idx = (int32_t*)(val + npow);            // [npow]
    uint32_t* cho = (uint32_t*)(idx + npow);          // [S] 4 bits per group
    for (int c = blockIdx.x; c < n_clusters; c += gridDim.x) {
        const int seg0 = cl_off[c], nseg = cl_off[c + 1] - seg0, ng = n_groups[c];
        for (int s = threadIdx.x; s < npow; s += blockDim.x) {
            float v = 3.0e38f;
            uint32_t choice = 0;
            if (s < S) {
                if (ng == 1) {
                    const int nc = seg_off[seg0 + 1] - seg_off[seg0];      // copies of the chromosome (those of the first segment, :199)
                    for (int k = 0; k < nc; k++) {
                        const float r = sprite_rg2_of(P, s, seg0, nseg, seg_off, beads, seg_group, (uint32_t)k);
                        if (r < v) { v = r; choice = (uint32_t)k; }
                    }
                } else {
                    // copy selection on the representatives (get_rg2s_cpp over ng regions)
                    int rs[SPRITE_MAX_GROUPS], ncopy[SPRITE_MAX_GROUPS], ncomb = 1;
#pragma unroll
                    for (int g = 0; g < SPRITE_MAX_GROUPS; g++) {
                        rs[g] = 0; ncopy[g] = 1;
                        if (g < ng) { rs[g] = seg0 + rep[c * SPRITE_MAX_GROUPS + g]; ncopy[g] = seg_off[rs[g] + 1] - seg_off[rs[g]]; ncomb *= ncopy[g]; }
                    }
                    const float fg = (float)ng;
                    float best = 100000000.0f;
                    int best_k = 0;
                    for (int k = 0; k < ncomb; k++) {
                        float px[SPRITE_MAX_GROUPS], py[SPRITE_MAX_GROUPS], pz[SPRITE_MAX_GROUPS];
                        float mx = 0.f, my = 0.f, mz = 0.f;
                        int kk = k;
#pragma unroll
                        for (int g = 0; g < SPRITE_MAX_GROUPS; g++) {
                            if (g < ng) {
                                const int si = kk % ncopy[g]; kk /= ncopy[g];
                                const float* x = P.crd + ((size_t)beads[seg_off[rs[g]] + si] * S + s) * 3;
                                px[g] = x[0]; py[g] = x[1]; pz[g] = x[2];
                                mx = __fadd_rn(mx, px[g]); my = __fadd_rn(my, py[g]); mz = __fadd_rn(mz, pz[g]);
                            }
                        }
                        mx = __fdiv_rn(mx, fg); my = __fdiv_rn(my, fg); mz = __fdiv_rn(mz, fg);
                        float rg = 0.f;
#pragma unroll
                        for (int g = 0; g < SPRITE_MAX_GROUPS; g++) {
                            if (g < ng) {
                                const float dx = __fsub_rn(px[g], mx), dy = __fsub_rn(py[g], my), dz = __fsub_rn(pz[g], mz);
                                rg = __fadd_rn(rg, __fadd_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)), __fmul_rn(dz, dz)));
                            }
                        }
                        rg = __fdiv_rn(rg, fg);
                        if (rg < best) { best = rg; best_k = k; }
                    }
                    int kk = best_k;
#pragma unroll
                    for (int g = 0; g < SPRITE_MAX_GROUPS; g++)
                        if (g < ng) { choice |= (uint32_t)(kk % ncopy[g]) << (4 * g); kk /= ncopy[g]; }
                    v = sprite_rg2_of(P, s, seg0, nseg, seg_off, beads, seg_group, choice);       // all segments, chosen copies (:262-283)
                }
                cho[s] = choice;
            }
            val[s] = v; idx[s] = s < S ? s : 0x7fffffff;
        }
        __syncthreads();
        // bitonic sort of (value, structure) ascending
        for (int k = 2; k <= npow; k <<= 1)
            for (int j = k >> 1; j > 0; j >>= 1) {
                for (int t = threadIdx.x; t < npow; t += blockDim.x) {
                    const int o = t ^ j;
                    if (o > t) {
                        const bool up = (t & k) == 0;
                        const float a = val[t], b = val[o];
                        const int ia = idx[t], ib = idx[o];
                        const bool gt = a > b || (a == b && ia > ib);
                        if (gt == up) { val[t] = b; val[o] = a; idx[t] = ib; idx[o] = ia; }
                    }
                }
                __syncthreads();
            }
        for (int j = threadIdx.x; j < keep; j += blockDim.x) {
            best_rg2[(size_t)c * keep + j] = val[j];
            best_struct[(size_t)c * keep + j] = idx[j];
        }
        for (int q = threadIdx.x; q < nseg * keep; q += blockDim.x) {
            const int i = q / keep, j = q % keep;
            const uint32_t choice = cho[idx[j]];
            best_beads[(size_t)(seg0 + i) * keep + j] = beads[seg_off[seg0 + i] + ((choice >> (4 * seg_group[seg0 + i])) & 15)];
        }
        __syncthreads();
    }
}

}  // namespace igmk