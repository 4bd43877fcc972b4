// K1: device-resident encoder (text -> 4-bit IUPAC sets as three bit-planes + a sparse list of ambiguity codes,
// column-tiled: see common.cuh) + reference filter counters,
// and the synthetic alignment generator used by bench.py.  Replaces, on the device,
//   quick_count_sequence_non_N            src/fastaseq.c:642-648   (non_n over the WHOLE row)
//   the ref filter                        src/nearest.c:237,263-270
//   char2bit lookups of the scoring loop  biomcmc alignment.c:868-870, 906-931
#include "common.cuh"
#include "kernels.h"

namespace uvg {

__constant__ uint8_t c_lut[3][256];    // [0] default-mode code, [1] --acgt-mode code, [2] valid text (not in "NnXx-?Oo.")

void build_luts(uint8_t lut[3][256])
{
    for (int i = 0; i < 256; ++i) { lut[0][i] = 0; lut[1][i] = 0; lut[2][i] = 1; }
    // IUPAC sets; N O X ? - and unknown bytes stay 0 (invalid site).  U == T in default mode (alignment.c:912-931).
    const char* sym = "ACGTUMRWSYKVHDB";
    const uint8_t val[] = {1, 2, 4, 8, 8, 3, 5, 9, 6, 10, 12, 7, 11, 13, 14};
    for (int k = 0; sym[k]; ++k) {
        lut[0][(uint8_t)sym[k]] = val[k];
        lut[0][(uint8_t)(sym[k] + 32)] = val[k];            // the reference upper-cases on read (fastaseq.c:451-452)
    }
    // --acgt mode: only A C G T are usable (is_acgt[], src/utils.c:262; U is NOT)
    const char* acgt = "ACGT";
    for (int k = 0; k < 4; ++k) { lut[1][(uint8_t)acgt[k]] = (uint8_t)(1 << k); lut[1][(uint8_t)(acgt[k] + 32)] = (uint8_t)(1 << k); }
    const char* indel = "NnXx-?Oo.";                         // is_indel[], src/utils.c:263
    for (int k = 0; indel[k]; ++k) lut[2][(uint8_t)indel[k]] = 0;
}

cudaError_t upload_luts()
{
    uint8_t lut[3][256];
    build_luts(lut);
    return cudaMemcpyToSymbol(c_lut, lut, sizeof(lut));
}

// ---- non_n + accept flag + number of ambiguity codes: one block per row ----------------------------------------------
__global__ void __launch_bounds__(256) k_count_non_n(const uint8_t* __restrict__ text, int64_t n, int nchar, int min_non_n, int acgt_mode,
                                                     int32_t* __restrict__ non_n, uint8_t* __restrict__ accepted, int32_t* __restrict__ n_amb)
{
    // s_tab[b] = valid-text flag (bit 0) | ambiguity-code flag << 16, so one shared-memory lookup per byte feeds both counters
    __shared__ uint32_t s_tab[256];
    __shared__ uint32_t s_part[8];
    {
        const uint32_t code = c_lut[acgt_mode ? 1 : 0][threadIdx.x];
        s_tab[threadIdx.x] = (uint32_t)c_lut[2][threadIdx.x] | ((code & (code - 1u)) ? 0x10000u : 0u);
    }
    __syncthreads();
    for (int64_t row = blockIdx.x; row < n; row += gridDim.x) {
        const uint8_t* p = text + row * (int64_t)nchar;
        uint32_t cnt = 0;
        // head bytes up to 16-byte alignment, then 128-bit loads, then the tail
        int head = (int)((16 - ((uintptr_t)p & 15)) & 15);
        if (head > nchar) head = nchar;
        if ((int)threadIdx.x < head) cnt += s_tab[p[threadIdx.x]];
        int nvec = (nchar - head) >> 4;
        const uint4* pv = reinterpret_cast<const uint4*>(p + head);
        for (int i = threadIdx.x; i < nvec; i += blockDim.x) {
            uint4 v = __ldg(pv + i);
            uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
            for (int k = 0; k < 4; ++k)
                cnt += s_tab[w[k] & 255] + s_tab[(w[k] >> 8) & 255] + s_tab[(w[k] >> 16) & 255] + s_tab[w[k] >> 24];
        }
        int tail0 = head + (nvec << 4);
        if (tail0 + (int)threadIdx.x < nchar) cnt += s_tab[p[tail0 + threadIdx.x]];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) cnt += __shfl_xor_sync(0xffffffffu, cnt, o);      // both 16-bit halves stay < 65536 per row chunk
        if ((threadIdx.x & 31) == 0) s_part[threadIdx.x >> 5] = cnt;
        __syncthreads();
        if (threadIdx.x == 0) {
            int tot = 0, amb = 0;
            for (int k = 0; k < 8; ++k) { tot += (int)(s_part[k] & 0xffffu); amb += (int)(s_part[k] >> 16); }
            non_n[row] = tot;
            accepted[row] = (uint8_t)(tot >= min_non_n);      // nearest.c:264
            n_amb[row] = amb;
        }
        __syncthreads();
    }
}

// long rows: the packed 16-bit halves above would wrap beyond 65,535 sites per thread-strided chunk -> wide variant
__global__ void __launch_bounds__(256) k_count_non_n_wide(const uint8_t* __restrict__ text, int64_t n, int nchar, int min_non_n, int acgt_mode,
                                                          int32_t* __restrict__ non_n, uint8_t* __restrict__ accepted, int32_t* __restrict__ n_amb)
{
    __shared__ uint8_t s_valid[256], s_amb[256];
    __shared__ int s_v[8], s_a[8];
    {
        const uint32_t code = c_lut[acgt_mode ? 1 : 0][threadIdx.x];
        s_valid[threadIdx.x] = c_lut[2][threadIdx.x];
        s_amb[threadIdx.x] = (code & (code - 1u)) ? 1 : 0;
    }
    __syncthreads();
    for (int64_t row = blockIdx.x; row < n; row += gridDim.x) {
        const uint8_t* p = text + row * (int64_t)nchar;
        int v = 0, a = 0;
        for (int i = threadIdx.x; i < nchar; i += blockDim.x) { const uint8_t b = __ldg(p + i); v += s_valid[b]; a += s_amb[b]; }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) { v += __shfl_xor_sync(0xffffffffu, v, o); a += __shfl_xor_sync(0xffffffffu, a, o); }
        if ((threadIdx.x & 31) == 0) { s_v[threadIdx.x >> 5] = v; s_a[threadIdx.x >> 5] = a; }
        __syncthreads();
        if (threadIdx.x == 0) {
            int tot = 0, amb = 0;
            for (int k = 0; k < 8; ++k) { tot += s_v[k]; amb += s_a[k]; }
            non_n[row] = tot;
            accepted[row] = (uint8_t)(tot >= min_non_n);
            n_amb[row] = amb;
        }
        __syncthreads();
    }
}

// ---- exclusive scans over the accepted rows: destination local index (or -1) and ambiguity-entry offset; single block ----
__global__ void __launch_bounds__(1024) k_scan_accept(const uint8_t* __restrict__ accepted, const int32_t* __restrict__ n_amb, int64_t n,
                                                      int64_t dst_base, uint32_t amb_base, int64_t* __restrict__ dst,
                                                      uint32_t* __restrict__ ramb_off, int64_t* __restrict__ totals)
{
    __shared__ int s_warp[32], s_warp2[32];
    __shared__ int64_t s_run, s_run2;
    if (threadIdx.x == 0) { s_run = 0; s_run2 = 0; }
    __syncthreads();
    for (int64_t base = 0; base < n; base += blockDim.x) {
        int64_t i = base + threadIdx.x;
        int a = (i < n) ? accepted[i] : 0;
        int m = (a && i < n) ? n_amb[i] : 0;
        int incl = a, incl2 = m;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            int t = __shfl_up_sync(0xffffffffu, incl, o), t2 = __shfl_up_sync(0xffffffffu, incl2, o);
            if ((int)(threadIdx.x & 31) >= o) { incl += t; incl2 += t2; }
        }
        if ((threadIdx.x & 31) == 31) { s_warp[threadIdx.x >> 5] = incl; s_warp2[threadIdx.x >> 5] = incl2; }
        __syncthreads();
        if (threadIdx.x < 32) {
            int v = s_warp[threadIdx.x], inc = v, v2 = s_warp2[threadIdx.x], inc2 = v2;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                int t = __shfl_up_sync(0xffffffffu, inc, o), t2 = __shfl_up_sync(0xffffffffu, inc2, o);
                if ((int)threadIdx.x >= o) { inc += t; inc2 += t2; }
            }
            s_warp[threadIdx.x] = inc - v;                    // exclusive prefix of warp totals
            s_warp2[threadIdx.x] = inc2 - v2;
        }
        __syncthreads();
        int excl = incl - a + s_warp[threadIdx.x >> 5];
        int excl2 = incl2 - m + s_warp2[threadIdx.x >> 5];
        int64_t run = s_run, run2 = s_run2;
        if (i < n) {
            dst[i] = a ? dst_base + run + excl : -1;
            if (a) ramb_off[dst_base + run + excl] = amb_base + (uint32_t)(run2 + excl2);
        }
        __syncthreads();
        if (threadIdx.x == blockDim.x - 1) { s_run = run + excl + a; s_run2 = run2 + excl2 + m; }
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        totals[0] = s_run; totals[1] = s_run2;
        ramb_off[dst_base + s_run] = amb_base + (uint32_t)s_run2;          // closing offset of the CSR
    }
}

// ---- encoder: one block per source row; a warp turns 128 sites into one uint4 unit per plane with ballots ----------
__global__ void __launch_bounds__(256) k_encode(const uint8_t* __restrict__ text, int64_t n, int nchar, int Wp, int acgt_mode,
                                                const int64_t* __restrict__ dst, const int32_t* __restrict__ non_n_src,
                                                uint4* __restrict__ db, const uint32_t* __restrict__ ramb_off, uint32_t* __restrict__ ramb_ent,
                                                int32_t* __restrict__ non_n_db)
{
    __shared__ uint8_t s_code[256];
    __shared__ uint32_t s_cur;
    s_code[threadIdx.x] = c_lut[acgt_mode ? 1 : 0][threadIdx.x];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarp = blockDim.x >> 5;
    for (int64_t row = blockIdx.x; row < n; row += gridDim.x) {
        const int64_t d = dst[row];
        __syncthreads();
        if (d < 0) continue;
        if (threadIdx.x == 0) { non_n_db[d] = non_n_src[row]; s_cur = ramb_off[d]; }
        __syncthreads();
        const uint8_t* p = text + row * (int64_t)nchar;
        const int64_t group = d / GROUP;
        const int rl = (int)(d % GROUP);
        for (int u = warp; u < Wp / 4; u += nwarp) {
            uint32_t pb[4], p0[4], p1[4];
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const int site = (u * 4 + k) * 32 + lane;
                const uint32_t c = (site < nchar) ? s_code[__ldg(p + site)] : 0u;
                const bool multi = (c & (c - 1u)) != 0u;
                const bool plain = c != 0u && !multi;
                pb[k] = __ballot_sync(0xffffffffu, plain);
                p0[k] = __ballot_sync(0xffffffffu, plain && (c & 10u));         // C or T
                p1[k] = __ballot_sync(0xffffffffu, plain && (c & 12u));         // G or T
                const uint32_t am = __ballot_sync(0xffffffffu, multi);
                if (am) {
                    uint32_t base = 0;
                    if (lane == 0) base = atomicAdd(&s_cur, (uint32_t)__popc(am));
                    base = __shfl_sync(0xffffffffu, base, 0);
                    if (multi) ramb_ent[base + __popc(am & ((1u << lane) - 1u))] = (uint32_t)site | (c << 28);
                }
            }
            if (lane < NPLANE) {
                const uint4 v = lane == 0 ? make_uint4(pb[0], pb[1], pb[2], pb[3])
                              : lane == 1 ? make_uint4(p0[0], p0[1], p0[2], p0[3]) : make_uint4(p1[0], p1[1], p1[2], p1[3]);
                db[db_unit_index(group, u, lane, rl, Wp)] = v;
            }
        }
    }
}

// ---- unpack hook (tests): planes + ambiguity entries back to 4-bit codes; one block per call -------------------------
__global__ void k_unpack_ref(const uint4* __restrict__ db, const uint32_t* __restrict__ ramb_off, const uint32_t* __restrict__ ramb_ent,
                             int64_t local, int nchar, int Wp, uint8_t* __restrict__ out)
{
    const int64_t group = local / GROUP;
    const int rl = (int)(local % GROUP);
    const uint32_t* words = reinterpret_cast<const uint32_t*>(db);
    for (int site = threadIdx.x; site < nchar; site += blockDim.x) {
        const int w = site >> 5, b = site & 31;
        uint32_t pl[NPLANE];
#pragma unroll
        for (int k = 0; k < NPLANE; ++k) pl[k] = words[db_unit_index(group, w >> 2, k, rl, Wp) * 4 + (w & 3)];
        const uint32_t base = (pl[0] >> b) & 1u, b0 = (pl[1] >> b) & 1u, b1 = (pl[2] >> b) & 1u;
        out[site] = (uint8_t)(base ? (1u << (b0 + 2u * b1)) : 0u);
    }
    __syncthreads();
    for (uint32_t e = ramb_off[local] + threadIdx.x; e < ramb_off[local + 1]; e += blockDim.x) {
        const uint32_t v = ramb_ent[e];
        out[v & AMB_SITE_MASK] = (uint8_t)(v >> 28);
    }
}

// ================================================================================================================
// synthetic rows (bench utility): same population model as uvaia_b200/synth.py, counter-based RNG
// ================================================================================================================
__device__ __forceinline__ uint64_t mix64(uint64_t z)
{
    z += 0x9e3779b97f4a7c15ull;
    z = (z ^ (z >> 30)) * 0xbf58476d1ce4e5b9ull;
    z = (z ^ (z >> 27)) * 0x94d049bb133111ebull;
    return z ^ (z >> 31);
}
struct Rng {
    uint64_t key, ctr;
    __device__ Rng(uint64_t seed, uint64_t ordinal) : key(mix64(seed ^ mix64(ordinal))), ctr(0) {}
    __device__ uint64_t next() { return mix64(key + (ctr++) * 0xd1342543de82ef95ull); }
    __device__ double unif() { return ((next() >> 11) + 0.5) * (1.0 / 9007199254740992.0); }
    __device__ int below(int n) { return (int)(next() % (uint64_t)n); }
    __device__ int poisson(double lam) {       // Knuth; lam is small (<= ~30)
        double L = exp(-lam), p = 1.0; int k = 0;
        do { ++k; p *= unif(); } while (p > L && k < 1000);
        return k - 1;
    }
    __device__ int geometric(double mean) {   // support 1,2,...
        if (mean <= 1.0) return 1;
        double u = unif();
        int v = 1 + (int)floor(log(u) / log(1.0 - 1.0 / mean));
        return v < 1 ? 1 : v;
    }
};

constexpr int SY_MAXSUB = 64, SY_MAXRUN = 96, SY_MAXGAP = 32, SY_MAXAMB = 16;

__global__ void __launch_bounds__(256) k_synth_rows(const uint8_t* __restrict__ clades, int n_clades, int nchar, uint64_t seed,
                                                    int64_t first, int64_t n, double n_frac_mean, double ambiguous_frac,
                                                    uint8_t* __restrict__ out)
{
    __shared__ int s_clade, s_nsub, s_nrun, s_ngap, s_namb, s_lead, s_trail;
    __shared__ int s_sub_pos[SY_MAXSUB], s_sub_alt[SY_MAXSUB];
    __shared__ int s_run_pos[SY_MAXRUN], s_run_len[SY_MAXRUN];
    __shared__ int s_gap_pos[SY_MAXGAP], s_gap_len[SY_MAXGAP];
    __shared__ int s_amb_pos[SY_MAXAMB], s_amb_code[SY_MAXAMB];
    for (int64_t row = blockIdx.x; row < n; row += gridDim.x) {
        const uint64_t ord = (uint64_t)(first + row);
        __syncthreads();
        if (threadIdx.x == 0) {
            Rng g(seed, ord);
            const double scale = nchar / 29903.0;
            s_clade = g.below(n_clades);
            int k = g.poisson(8.0); if (k > SY_MAXSUB) k = SY_MAXSUB;
            for (int i = 0; i < k; ++i) { s_sub_pos[i] = g.below(nchar); s_sub_alt[i] = g.below(3); }
            s_nsub = k;
            if (nchar > 400) { s_lead = (int)((30 + g.below(121)) * scale); s_trail = (int)((30 + g.below(91)) * scale); }
            else { s_lead = g.below(3); s_trail = g.below(3); }
            // per-genome N fraction: exponential with the requested mean (clipped), spread over ~Poisson(20) runs
            double f = -log(g.unif()) * n_frac_mean; if (f > 0.9) f = 0.9;
            int nr = g.poisson(n_frac_mean < 0.05 ? 2.0 : 20.0); if (nr > SY_MAXRUN) nr = SY_MAXRUN;
            double mean_len = nr ? fmax(1.0, nchar * f / nr) : 1.0;
            for (int i = 0; i < nr; ++i) { s_run_pos[i] = g.below(nchar); s_run_len[i] = g.geometric(mean_len); }
            s_nrun = nr;
            int ng = g.poisson(fmax(0.2, 0.0008 * nchar / 5.0)); if (ng > SY_MAXGAP) ng = SY_MAXGAP;
            for (int i = 0; i < ng; ++i) { s_gap_pos[i] = g.below(nchar); s_gap_len[i] = 1 + g.below(9); }
            s_ngap = ng;
            int na = g.poisson(1.7); if (na > SY_MAXAMB) na = SY_MAXAMB;
            const char two[] = "MRWSYK";
            for (int i = 0; i < na; ++i) { s_amb_pos[i] = g.below(nchar); s_amb_code[i] = two[g.below(6)]; }
            s_namb = na;
        }
        __syncthreads();
        uint8_t* o = out + row * (int64_t)nchar;
        const uint8_t* c = clades + (int64_t)s_clade * nchar;
        const uint64_t site_key = mix64(seed * 0x2545f4914f6cdd1dull + ord);
        for (int i = threadIdx.x; i < nchar; i += blockDim.x) {
            uint8_t b = c[i];
            if (i < s_lead || i >= nchar - s_trail) b = 'N';
            o[i] = b;
        }
        __syncthreads();
        if ((int)threadIdx.x < s_nsub) {        // private substitutions (applied before masking like synth.py)
            int pos = s_sub_pos[threadIdx.x];
            uint8_t cur = c[pos];
            const char* alt = cur == 'A' ? "CGT" : cur == 'C' ? "AGT" : cur == 'G' ? "ACT" : "ACG";
            if (pos >= s_lead && pos < nchar - s_trail) o[pos] = (uint8_t)alt[s_sub_alt[threadIdx.x]];
        }
        __syncthreads();
        for (int r = 0; r < s_nrun; ++r) {
            int p0 = s_run_pos[r], len = s_run_len[r];
            for (int i = threadIdx.x; i < len && p0 + i < nchar; i += blockDim.x) o[p0 + i] = 'N';
        }
        __syncthreads();
        for (int r = 0; r < s_ngap; ++r) {
            int p0 = s_gap_pos[r], len = s_gap_len[r];
            if ((int)threadIdx.x < len && p0 + (int)threadIdx.x < nchar) o[p0 + threadIdx.x] = '-';
        }
        __syncthreads();
        if ((int)threadIdx.x < s_namb) o[s_amb_pos[threadIdx.x]] = (uint8_t)s_amb_code[threadIdx.x];
        __syncthreads();
        if (ambiguous_frac > 0.0) {              // config-4 stress: 3/4 N, 1/4 IUPAC (two- and three-base codes)
            const char codes[] = "MRWSYKVHDB";
            for (int i = threadIdx.x; i < nchar; i += blockDim.x) {
                uint64_t h = mix64(site_key + (uint64_t)i * 0x9e3779b97f4a7c15ull);
                double u = (h >> 11) * (1.0 / 9007199254740992.0);
                if (u < ambiguous_frac) o[i] = (u < 0.75 * ambiguous_frac) ? (uint8_t)'N' : (uint8_t)codes[(h & 1023) % 10];
            }
        }
    }
}

// ---- host launchers ------------------------------------------------------------------------------------------------
cudaError_t launch_count_non_n(const uint8_t* d_text, int64_t n, int nchar, int min_non_n, int acgt_mode, int32_t* d_non_n, uint8_t* d_acc,
                               int32_t* d_namb, cudaStream_t st)
{
    if (n <= 0) return cudaSuccess;
    int grid = (int)(n < 148 * 32 ? n : 148 * 32);
    if (nchar < 65536) k_count_non_n<<<grid, 256, 0, st>>>(d_text, n, nchar, min_non_n, acgt_mode, d_non_n, d_acc, d_namb);
    else k_count_non_n_wide<<<grid, 256, 0, st>>>(d_text, n, nchar, min_non_n, acgt_mode, d_non_n, d_acc, d_namb);
    return cudaGetLastError();
}
cudaError_t launch_scan_accept(const uint8_t* d_acc, const int32_t* d_namb, int64_t n, int64_t dst_base, uint32_t amb_base, int64_t* d_dst,
                               uint32_t* ramb_off, int64_t* d_totals, cudaStream_t st)
{
    k_scan_accept<<<1, 1024, 0, st>>>(d_acc, d_namb, n, dst_base, amb_base, d_dst, ramb_off, d_totals);
    return cudaGetLastError();
}
cudaError_t launch_encode(const uint8_t* d_text, int64_t n, int nchar, int Wp, int acgt_mode, const int64_t* d_dst,
                          const int32_t* d_non_n_src, uint4* db, const uint32_t* ramb_off, uint32_t* ramb_ent, int32_t* non_n_db, cudaStream_t st)
{
    if (n <= 0) return cudaSuccess;
    int grid = (int)(n < 148 * 16 ? n : 148 * 16);
    k_encode<<<grid, 256, 0, st>>>(d_text, n, nchar, Wp, acgt_mode, d_dst, d_non_n_src, db, ramb_off, ramb_ent, non_n_db);
    return cudaGetLastError();
}
cudaError_t launch_unpack_ref(const uint4* db, const uint32_t* ramb_off, const uint32_t* ramb_ent, int64_t local, int nchar, int Wp,
                              uint8_t* d_out, cudaStream_t st)
{
    k_unpack_ref<<<1, 256, 0, st>>>(db, ramb_off, ramb_ent, local, nchar, Wp, d_out);
    return cudaGetLastError();
}
cudaError_t launch_synth_rows(const uint8_t* d_clades, int n_clades, int nchar, uint64_t seed, int64_t first, int64_t n,
                              double n_frac_mean, double ambiguous_frac, uint8_t* d_out, cudaStream_t st)
{
    if (n <= 0) return cudaSuccess;
    int grid = (int)(n < 148 * 16 ? n : 148 * 16);
    k_synth_rows<<<grid, 256, 0, st>>>(d_clades, n_clades, nchar, seed, first, n, n_frac_mean, ambiguous_frac, d_out);
    return cudaGetLastError();
}

}  // namespace uvg
