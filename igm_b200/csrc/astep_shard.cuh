// astep_shard.cuh -- Hi-C activation distances when the population stays SHARDED by structure over the GPUs
// (north_star: "for the A-step NCCL over NVLink ... reduces the per-pair distance histograms").
//
// get_actdist (igm/steps/ActivationDistanceStep.py:319-459) needs, per haploid pair, two population-wide numbers:
// the count of squared distances <= rcutsq (:419) and the o-th smallest squared distance (:425,:446).  Both are
// sums over structures, so every rank computes them for ITS structures and NCCL all-reduces (sum):
//   k_shard_count   local contact count per pair                       -> all-reduce -> pnow, p, o   (k_shard_rank)
//   k_shard_hist    local 256-bin histogram of one radix digit of the float32 bit patterns that match the prefix
//                   found so far                                        -> all-reduce -> next digit  (k_shard_pick)
// four digits (most significant first) pin the exact bit pattern of the o-th smallest value -- bit-identical to
// the single-GPU kernels whatever the number of ranks.  Arithmetic as in astep.cuh (float32, no FMA contraction).
#pragma once
#include "astep.cuh"

namespace igmk {

// every squared distance of pair `pidx` in this rank's structures goes through `fn(float v)`
template <typename F>
__device__ __forceinline__ void shard_for_each_dsq(const PopView& P, int hi, int hj, int tid, int nthreads, F fn)
{
    const int S = P.n_struct;
    const int io = P.copy_off[hi], jo = P.copy_off[hj];
    const int ni = P.copy_off[hi + 1] - io, nj = P.copy_off[hj + 1] - jo;
    const bool intra = P.chrom[hi] == P.chrom[hj];
    const int ncomb = intra ? min(ni, nj) : ni * nj;          // ActivationDistanceStep.py:384-385, :400-401
    for (int c = 0; c < ncomb; c++) {
        const int bi = P.copy_beads[io + (intra ? c : c / nj)];
        const int bj = P.copy_beads[jo + (intra ? c : c % nj)];
        const float* xi = P.crd + (size_t)bi * S * 3;
        const float* xj = P.crd + (size_t)bj * S * 3;
        for (int s = tid; s < S; s += nthreads) {
            const float dx = __fsub_rn(xi[3 * s], xj[3 * s]);
            const float dy = __fsub_rn(xi[3 * s + 1], xj[3 * s + 1]);
            const float dz = __fsub_rn(xi[3 * s + 2], xj[3 * s + 2]);
            fn(__fadd_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)), __fmul_rn(dz, dz)));
        }
    }
}

// local contact counts, one CTA per pair
__global__ void __launch_bounds__(AD_THREADS)
k_shard_count(PopView P, int64_t n_pairs, const int32_t* __restrict__ pair_i, const int32_t* __restrict__ pair_j,
              float contact_range_f32, uint32_t* __restrict__ counts)
{
    __shared__ int s_red[AD_THREADS / 32];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    for (int64_t pidx = blockIdx.x; pidx < n_pairs; pidx += gridDim.x) {
        const int hi = pair_i[pidx], hj = pair_j[pidx];
        int cnt = 0;
        if (hi != hj) {
            const float ri = P.radii[P.copy_beads[P.copy_off[hi]]], rj = P.radii[P.copy_beads[P.copy_off[hj]]];
            const float rc = __fmul_rn(contact_range_f32, __fadd_rn(ri, rj));
            const float rcutsq = __fmul_rn(rc, rc);           // :368
            shard_for_each_dsq(P, hi, hj, tid, AD_THREADS, [&](float v) { cnt += (v <= rcutsq); });
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) cnt += __shfl_xor_sync(0xffffffffu, cnt, o);
        if (lane == 0) s_red[warp] = cnt;
        __syncthreads();
        if (tid == 0) {
            int t = 0;
            for (int w = 0; w < AD_THREADS / 32; w++) t += s_red[w];
            counts[pidx] = (uint32_t)t;
        }
        __syncthreads();
    }
}

// population-wide counts -> corrected probability and the wanted rank o (every rank computes the same numbers)
__global__ void __launch_bounds__(256)
k_shard_rank(PopView P, int64_t n_pairs, int32_t n_struct_total, const int32_t* __restrict__ pair_i, const int32_t* __restrict__ pair_j,
             const double* __restrict__ pwish, const double* __restrict__ plast, const uint32_t* __restrict__ counts_total,
             uint32_t* __restrict__ want, uint32_t* __restrict__ prefix, float* __restrict__ out_prob, uint8_t* __restrict__ out_valid)
{
    for (int64_t pidx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; pidx < n_pairs; pidx += (int64_t)gridDim.x * blockDim.x) {
        const int hi = pair_i[pidx], hj = pair_j[pidx];
        prefix[pidx] = 0u; want[pidx] = 0u;
        if (hi == hj) { out_valid[pidx] = 0; continue; }      // :351-352
        const int ni = P.copy_off[hi + 1] - P.copy_off[hi], nj = P.copy_off[hj + 1] - P.copy_off[hj];
        const int ncomb = P.chrom[hi] == P.chrom[hj] ? min(ni, nj) : ni * nj;
        const long long total = (long long)ncomb * n_struct_total;
        const double pnow = (double)counts_total[pidx] / (double)total;                  // :422
        const double pl = plast[pidx], pw = pwish[pidx];
        double t = pl < 1.0 ? (pnow - pl) / (1.0 - pl) : pnow; t = t > 0.0 ? t : 0.0;    // :430
        double p = t < 1.0 ? (pw - t) / (1.0 - t) : pw; p = p > 0.0 ? p : 0.0;           // :434
        if (!(p > 0.0)) { out_valid[pidx] = 0; continue; }
        const double prod = ((double)ncomb * p) * (double)n_struct_total;                // :443-444
        long long o = (long long)rint(prod);
        if (o > total - 1) o = total - 1;
        want[pidx] = (uint32_t)o;
        out_prob[pidx] = text_roundtrip4(p);
        out_valid[pidx] = 1;
    }
}

// local histogram of radix digit `pass` (0 = bits 31..24) of the keys whose higher digits equal prefix[pidx]
__global__ void __launch_bounds__(AD_THREADS)
k_shard_hist(PopView P, int64_t n_pairs, const int32_t* __restrict__ pair_i, const int32_t* __restrict__ pair_j,
             const uint8_t* __restrict__ valid, const uint32_t* __restrict__ prefix, int pass, uint32_t* __restrict__ hist)
{
    __shared__ uint32_t s_hist[256];
    const int tid = threadIdx.x;
    const int shift = 24 - 8 * pass;
    const uint32_t hmask = pass == 0 ? 0u : (0xffffffffu << (shift + 8));
    for (int64_t pidx = blockIdx.x; pidx < n_pairs; pidx += gridDim.x) {
        s_hist[tid] = 0u;
        __syncthreads();
        if (valid[pidx]) {
            const uint32_t pre = prefix[pidx];
            shard_for_each_dsq(P, pair_i[pidx], pair_j[pidx], tid, AD_THREADS, [&](float v) {
                const uint32_t key = fkey(v);
                if ((key & hmask) == pre) atomicAdd(&s_hist[(key >> shift) & 0xffu], 1u);
            });
        }
        __syncthreads();
        hist[pidx * 256 + tid] = s_hist[tid];
        __syncthreads();
    }
}

// population-wide histogram -> the digit that holds rank want[pidx]; after the last pass the activation distance
__global__ void __launch_bounds__(256)
k_shard_pick(int64_t n_pairs, const uint8_t* __restrict__ valid, const uint32_t* __restrict__ hist_total, int pass,
             uint32_t* __restrict__ want, uint32_t* __restrict__ prefix, float* __restrict__ out_dist)
{
    const int lane = threadIdx.x & 31;
    const int64_t warp_global = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int64_t n_warps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    const int shift = 24 - 8 * pass;
    for (int64_t pidx = warp_global; pidx < n_pairs; pidx += n_warps) {
        if (!valid[pidx]) continue;
        const uint32_t* h = hist_total + pidx * 256;
        uint32_t k = want[pidx];
        // lane l owns bins 8l .. 8l+7
        uint32_t v[8], sum = 0;
#pragma unroll
        for (int b = 0; b < 8; b++) { v[b] = h[8 * lane + b]; sum += v[b]; }
        uint32_t incl = sum;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { const uint32_t t = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += t; }
        const uint32_t excl = incl - sum;
        const bool mine = k >= excl && k < incl;              // exactly one lane: k < total live keys
        uint32_t digit = 0, rest = 0;
        if (mine) {
            uint32_t run = excl;
#pragma unroll
            for (int b = 0; b < 8; b++) {
                if (k >= run && k < run + v[b]) { digit = 8u * lane + b; rest = k - run; }
                run += v[b];
            }
        }
        const uint32_t who = __ballot_sync(0xffffffffu, mine);
        if (!who) continue;                                    // (cannot happen: want < number of live keys)
        const int src = __ffs(who) - 1;
        digit = __shfl_sync(0xffffffffu, digit, src);
        rest = __shfl_sync(0xffffffffu, rest, src);
        if (lane == 0) {
            const uint32_t pre = prefix[pidx] | (digit << shift);
            prefix[pidx] = pre;
            want[pidx] = rest;
            if (pass == 3) out_dist[pidx] = text_roundtrip4(sqrt((double)__uint_as_float(pre)));   // :446 + the text round trip
        }
    }
}

}  // namespace igmk
