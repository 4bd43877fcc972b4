// K1: device-resident encoder (text -> 4-bit IUPAC bit-planes, column-tiled) + reference filter counters,
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

// ---- non_n + accept flag: one block per row -------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_count_non_n(const uint8_t* __restrict__ text, int64_t n, int nchar, int min_non_n,
                                                     int32_t* __restrict__ non_n, uint8_t* __restrict__ accepted)
{
    __shared__ uint8_t s_valid[256];
    __shared__ int s_part[8];
    s_valid[threadIdx.x] = c_lut[2][threadIdx.x];
    __syncthreads();
    for (int64_t row = blockIdx.x; row < n; row += gridDim.x) {
        const uint8_t* p = text + row * (int64_t)nchar;
        int cnt = 0;
        // head bytes up to 16-byte alignment, then 128-bit loads, then the tail
        int head = (int)((16 - ((uintptr_t)p & 15)) & 15);
        if (head > nchar) head = nchar;
        if ((int)threadIdx.x < head) cnt += s_valid[p[threadIdx.x]];
        int nvec = (nchar - head) >> 4;
        const uint4* pv = reinterpret_cast<const uint4*>(p + head);
        for (int i = threadIdx.x; i < nvec; i += blockDim.x) {
            uint4 v = __ldg(pv + i);
            uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
            for (int k = 0; k < 4; ++k)
                cnt += s_valid[w[k] & 255] + s_valid[(w[k] >> 8) & 255] + s_valid[(w[k] >> 16) & 255] + s_valid[w[k] >> 24];
        }
        int tail0 = head + (nvec << 4);
        if (tail0 + (int)threadIdx.x < nchar) cnt += s_valid[p[tail0 + threadIdx.x]];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) cnt += __shfl_xor_sync(0xffffffffu, cnt, o);
        if ((threadIdx.x & 31) == 0) s_part[threadIdx.x >> 5] = cnt;
        __syncthreads();
        if (threadIdx.x == 0) {
            int tot = 0;
            for (int k = 0; k < 8; ++k) tot += s_part[k];
            non_n[row] = tot;
            accepted[row] = (uint8_t)(tot >= min_non_n);      // nearest.c:264
        }
        __syncthreads();
    }
}

// ---- exclusive scan of the accept flags -> destination local index (or -1); single block ----------------------
__global__ void __launch_bounds__(1024) k_scan_accept(const uint8_t* __restrict__ accepted, int64_t n, int64_t dst_base,
                                                      int64_t* __restrict__ dst, int64_t* __restrict__ n_accepted)
{
    __shared__ int s_warp[32];
    __shared__ int64_t s_run;
    if (threadIdx.x == 0) s_run = 0;
    __syncthreads();
    for (int64_t base = 0; base < n; base += blockDim.x) {
        int64_t i = base + threadIdx.x;
        int a = (i < n) ? accepted[i] : 0;
        int incl = a;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { int t = __shfl_up_sync(0xffffffffu, incl, o); if ((int)(threadIdx.x & 31) >= o) incl += t; }
        if ((threadIdx.x & 31) == 31) s_warp[threadIdx.x >> 5] = incl;
        __syncthreads();
        if (threadIdx.x < 32) {
            int v = s_warp[threadIdx.x], inc2 = v;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) { int t = __shfl_up_sync(0xffffffffu, inc2, o); if ((int)threadIdx.x >= o) inc2 += t; }
            s_warp[threadIdx.x] = inc2 - v;                    // exclusive prefix of warp totals
        }
        __syncthreads();
        int excl = incl - a + s_warp[threadIdx.x >> 5];
        int64_t run = s_run;
        if (i < n) dst[i] = a ? dst_base + run + excl : -1;
        __syncthreads();
        if (threadIdx.x == blockDim.x - 1) s_run = run + excl + a;
        __syncthreads();
    }
    if (threadIdx.x == 0) *n_accepted = s_run;
}

// ---- encoder: one block per source row; warp-ballot transposes 32 bytes into four plane words ------------------
__global__ void __launch_bounds__(256) k_encode(const uint8_t* __restrict__ text, int64_t n, int nchar, int W, int Wp, int ambw,
                                                int acgt_mode, const int64_t* __restrict__ dst, const int32_t* __restrict__ non_n_src,
                                                uint4* __restrict__ db, uint32_t* __restrict__ refamb, int32_t* __restrict__ non_n_db)
{
    __shared__ uint8_t s_code[256];
    s_code[threadIdx.x] = c_lut[acgt_mode ? 1 : 0][threadIdx.x];
    __syncthreads();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarp = blockDim.x >> 5;
    for (int64_t row = blockIdx.x; row < n; row += gridDim.x) {
        const int64_t d = dst[row];
        if (d < 0) continue;
        const uint8_t* p = text + row * (int64_t)nchar;
        const int64_t group = d / GROUP;
        const int rl = (int)(d % GROUP);
        if (threadIdx.x == 0) non_n_db[d] = non_n_src[row];
        for (int w = warp; w < Wp; w += nwarp) {
            int site = w * 32 + lane;
            uint32_t c = (w < W && site < nchar) ? s_code[__ldg(p + site)] : 0u;
            uint32_t a = __ballot_sync(0xffffffffu, c & 1u);
            uint32_t cc = __ballot_sync(0xffffffffu, c & 2u);
            uint32_t g = __ballot_sync(0xffffffffu, c & 4u);
            uint32_t t = __ballot_sync(0xffffffffu, c & 8u);
            uint32_t amb = __ballot_sync(0xffffffffu, (c & (c - 1u)) != 0u);
            if (lane == 0) {
                db[db_word_index(group, w, rl, Wp)] = make_uint4(a, cc, g, t);
                if (amb) atomicOr(refamb + group * ambw + (w >> 5), 1u << (w & 31));
            }
        }
    }
}

// ---- unpack hook (tests): one block per call ---------------------------------------------------------------------
__global__ void k_unpack_ref(const uint4* __restrict__ db, int64_t local, int nchar, int Wp, uint8_t* __restrict__ out)
{
    const int64_t group = local / GROUP;
    const int rl = (int)(local % GROUP);
    for (int site = blockIdx.x * blockDim.x + threadIdx.x; site < nchar; site += gridDim.x * blockDim.x) {
        uint4 v = db[db_word_index(group, site >> 5, rl, Wp)];
        int b = site & 31;
        out[site] = (uint8_t)(((v.x >> b) & 1u) | (((v.y >> b) & 1u) << 1) | (((v.z >> b) & 1u) << 2) | (((v.w >> b) & 1u) << 3));
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
cudaError_t launch_count_non_n(const uint8_t* d_text, int64_t n, int nchar, int min_non_n, int32_t* d_non_n, uint8_t* d_acc,
                               cudaStream_t st)
{
    if (n <= 0) return cudaSuccess;
    int grid = (int)(n < 148 * 32 ? n : 148 * 32);
    k_count_non_n<<<grid, 256, 0, st>>>(d_text, n, nchar, min_non_n, d_non_n, d_acc);
    return cudaGetLastError();
}
cudaError_t launch_scan_accept(const uint8_t* d_acc, int64_t n, int64_t dst_base, int64_t* d_dst, int64_t* d_nacc, cudaStream_t st)
{
    k_scan_accept<<<1, 1024, 0, st>>>(d_acc, n, dst_base, d_dst, d_nacc);
    return cudaGetLastError();
}
cudaError_t launch_encode(const uint8_t* d_text, int64_t n, int nchar, int W, int Wp, int ambw, int acgt_mode, const int64_t* d_dst,
                          const int32_t* d_non_n_src, uint4* db, uint32_t* refamb, int32_t* non_n_db, cudaStream_t st)
{
    if (n <= 0) return cudaSuccess;
    int grid = (int)(n < 148 * 16 ? n : 148 * 16);
    k_encode<<<grid, 256, 0, st>>>(d_text, n, nchar, W, Wp, ambw, acgt_mode, d_dst, d_non_n_src, db, refamb, non_n_db);
    return cudaGetLastError();
}
cudaError_t launch_unpack_ref(const uint4* db, int64_t local, int nchar, int Wp, uint8_t* d_out, cudaStream_t st)
{
    k_unpack_ref<<<32, 256, 0, st>>>(db, local, nchar, Wp, d_out);
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
