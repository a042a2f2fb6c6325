// Stage 4 of the path: time-of-flight peak extraction on the upsampled frame.
//
// Replaces utils/mask2samples.py of the reference (paths relative to /root/reference):
//   nms_1d (:5-11)           max_pool1d(window, stride 1, -inf padding) + equality mask
//   thresholding (:14-23)    float threshold, or "keep entries equal to the row max"
//   get_maxima_positions (:26-34)  torch.nonzero of the result
//   mask2coords (:81-114)    bincount, K_max, ragged -> dense pack, divide by upsample_factor
//
// One CTA per row.  The row streams through shared memory in 4096-sample chunks (+ halo); each
// thread owns 16 consecutive samples in registers, builds the 21-wide sliding maximum from
// three levels of 3-input max (window 3 -> 9 -> 21, 78 max-ops per 16 samples), and the
// block compacts survivors in ascending order with warp-shuffle prefix sums.  The kernel is
// HBM-bound: algorithmic traffic is the 4*M bytes of the row, read once, plus the tiny coords table.
// Row-max mode (threshold None / 0.0): the maximum of the suppressed scores IS the row maximum whenever that is positive
// (the global maximum is always a window maximum, every other suppressed score is a smaller sample or 0), so the kept set is
// "every position equal to the row maximum" and one streaming arg-max-with-ties pass answers it without any window
// arithmetic (pk_rowmax_fast).  Rows with NaN / -inf entries, a non-positive maximum or too many ties take the literal
// two-pass route (nms -> row max of the suppressed scores -> selection), the second pass from L2.
//
// All comparisons follow the reference literally, including its NaN/inf behaviour:
//   sup = (x == wmax ? 1 : 0) * x     (0 * inf = NaN, and NaN counts as "non-zero")
//   wmax propagates NaN like ATen's max_pool1d; torch.max(dim) propagates NaN as well.
#include "common.cuh"

namespace stof {

constexpr int PK_THREADS = 256;
constexpr int PK_E = 16;                       // samples per thread per chunk
constexpr int PK_CHUNK = PK_THREADS * PK_E;    // 4096
constexpr int PK_HALO4 = 3;                    // left halo in float4 units (12 >= 10 samples)
constexpr int PK_Q4 = PK_CHUNK / 4 + 2 * PK_HALO4;          // 1030 logical float4 slots
constexpr int PK_SMEM4 = PK_Q4 + (PK_Q4 >> 2) + 1;           // padded: one spare float4 per 4
constexpr int PK_GEN_MAXH = 2048;              // generic-window kernel: max half window

__device__ __forceinline__ float max3nan(float a, float b, float c) {
    float d;
    asm("max.NaN.f32 %0, %1, %2, %3;" : "=f"(d) : "f"(a), "f"(b), "f"(c));
    return d;
}
__device__ __forceinline__ float max2nan(float a, float b) {
    float d;
    asm("max.NaN.f32 %0, %1, %2;" : "=f"(d) : "f"(a), "f"(b));
    return d;
}

// physical float4 slot of logical float4 slot q (bank-conflict-free for a lane stride of 4 slots)
__device__ __forceinline__ int pk_phys(int q) { return q + (q >> 2); }

// Stage row[cs-12 .. cs+4096+12) into shared memory; outside [0,M) reads as -inf
// (max_pool1d's implicit padding, utils/mask2samples.py:8).
// Unaligned rows: a global -> register load and a register -> shared store, so that the NEXT chunk's loads are in flight
// while the current chunk is processed.
constexpr int PK_NBUF = 2;                     // chunk buffers of the cp.async path (3 measured slower: 4.05 vs 4.17 TB/s, fewer CTAs per SM)
constexpr int PK_PRE = (PK_Q4 + PK_THREADS - 1) / PK_THREADS;     // float4 per thread per chunk (5)
__device__ __forceinline__ void pk_load(float4 (&pre)[PK_PRE], const float *__restrict__ row, int M, int cs) {
    const float ninf = -INFINITY;
#pragma unroll
    for (int k = 0; k < PK_PRE; ++k) {
        const int q = threadIdx.x + k * PK_THREADS;
        const int g = cs - 4 * PK_HALO4 + 4 * q;
        float4 v = make_float4(ninf, ninf, ninf, ninf);
        if (q < PK_Q4) {
            v.x = (g + 0 >= 0 && g + 0 < M) ? __ldg(row + g + 0) : ninf;
            v.y = (g + 1 >= 0 && g + 1 < M) ? __ldg(row + g + 1) : ninf;
            v.z = (g + 2 >= 0 && g + 2 < M) ? __ldg(row + g + 2) : ninf;
            v.w = (g + 3 >= 0 && g + 3 < M) ? __ldg(row + g + 3) : ninf;
        }
        pre[k] = v;
    }
}
__device__ __forceinline__ void pk_store(float4 *sm, const float4 (&pre)[PK_PRE]) {
#pragma unroll
    for (int k = 0; k < PK_PRE; ++k) {
        const int q = threadIdx.x + k * PK_THREADS;
        if (q < PK_Q4) sm[pk_phys(q)] = pre[k];
    }
}
// Aligned rows: the chunk goes global -> shared with 16-byte cp.async into the other of two chunk buffers (no registers held
// across the chunk's work, so four CTAs per SM each keep a whole 16 KB chunk in flight); float4 slots that straddle the row
// ends are filled with plain loads.
__device__ __forceinline__ void pk_fill_async(float4 *sm, const float *__restrict__ row, int M, int cs) {
    const float ninf = -INFINITY;
    if (cs >= 4 * PK_HALO4 && cs + PK_CHUNK + 4 * PK_HALO4 <= M) {
        // interior chunk: every slot lies inside the row, no per-slot bounds checks
        const float *src = row + cs - 4 * PK_HALO4;
#pragma unroll
        for (int k = 0; k < PK_Q4 / PK_THREADS; ++k) {
            const int q = threadIdx.x + k * PK_THREADS;
            const uint32_t d = (uint32_t)__cvta_generic_to_shared(sm + pk_phys(q));
            asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(d), "l"(src + 4 * q) : "memory");
        }
        if (threadIdx.x < PK_Q4 % PK_THREADS) {
            const int q = threadIdx.x + (PK_Q4 / PK_THREADS) * PK_THREADS;
            const uint32_t d = (uint32_t)__cvta_generic_to_shared(sm + pk_phys(q));
            asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(d), "l"(src + 4 * q) : "memory");
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
        return;
    }
    for (int q = threadIdx.x; q < PK_Q4; q += PK_THREADS) {
        const int g = cs - 4 * PK_HALO4 + 4 * q;
        float4 *dst = sm + pk_phys(q);
        if (g >= 0 && g + 3 < M) {
            const uint32_t d = (uint32_t)__cvta_generic_to_shared(dst);
            asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(d), "l"(row + g) : "memory");
        } else {
            float4 v;
            v.x = (g + 0 >= 0 && g + 0 < M) ? __ldg(row + g + 0) : ninf;
            v.y = (g + 1 >= 0 && g + 1 < M) ? __ldg(row + g + 1) : ninf;
            v.z = (g + 2 >= 0 && g + 2 < M) ? __ldg(row + g + 2) : ninf;
            v.w = (g + 3 >= 0 && g + 3 < M) ? __ldg(row + g + 3) : ninf;
            *dst = v;
        }
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
}

// sup[j] = (x == window21max ? 1 : 0) * x for the 16 samples owned by this thread
__device__ __forceinline__ void pk_nms21(const float4 *sm, float (&x)[PK_E], float (&sup)[PK_E]) {
    float v[40];
    const int q0 = 4 * threadIdx.x;
#pragma unroll
    for (int m = 0; m < 10; ++m) {
        const float4 t = sm[pk_phys(q0 + m)];
        v[4 * m + 0] = t.x; v[4 * m + 1] = t.y; v[4 * m + 2] = t.z; v[4 * m + 3] = t.w;
    }
    float a3[36];
#pragma unroll
    for (int i = 2; i <= 35; ++i) a3[i] = max3nan(v[i], v[i + 1], v[i + 2]);
    float a9[30];
#pragma unroll
    for (int i = 2; i <= 29; ++i) a9[i] = max3nan(a3[i], a3[i + 3], a3[i + 6]);
#pragma unroll
    for (int j = 0; j < PK_E; ++j) {
        const float wm = max3nan(a9[j + 2], a9[j + 8], a9[j + 14]);   // samples j+2 .. j+22 = x[j] +- 10
        x[j] = v[12 + j];
        sup[j] = (x[j] == wm ? 1.0f : 0.0f) * x[j];
    }
}

// generic odd window 2h+1 (any window_size of the reference API); chunk staged linearly
__device__ __forceinline__ float pk_nms_generic(const float *sm, int k, int h) {
    float wm = -INFINITY;
    for (int d = -h; d <= h; ++d) wm = max2nan(wm, sm[k + d]);
    const float x = sm[k];
    return (x == wm ? 1.0f : 0.0f) * x;
}

struct RowMax {
    float mx;
    int nan;
};

__device__ __forceinline__ RowMax block_rowmax(float mx, int nan, float *s_red) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
        nan |= __shfl_xor_sync(0xffffffffu, nan, o);
    }
    const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
    __syncthreads();
    if (l == 0) { s_red[w] = mx; s_red[8 + w] = __int_as_float(nan); }
    __syncthreads();
    mx = s_red[0]; nan = __float_as_int(s_red[8]);
#pragma unroll
    for (int i = 1; i < PK_THREADS / 32; ++i) { mx = fmaxf(mx, s_red[i]); nan |= __float_as_int(s_red[8 + i]); }
    RowMax r; r.mx = mx; r.nan = nan;
    return r;
}

// WIN21: specialised register path for window 21 (nms_win_size 20, config.yaml:26); otherwise generic.
// max.NaN / min over three inputs (FMNMX3 on sm_100): building blocks of the pre-filter in peaks_kernel
__device__ __forceinline__ float pk_max3nan(float a, float b, float c) {
    float r;
    asm("max.NaN.f32 %0, %1, %2, %3;" : "=f"(r) : "f"(a), "f"(b), "f"(c));
    return r;
}
__device__ __forceinline__ float pk_min3(float a, float b, float c) {
    float r;
    asm("min.f32 %0, %1, %2, %3;" : "=f"(r) : "f"(a), "f"(b), "f"(c));
    return r;
}


// ---- row-max mode, one streaming pass -----------------------------------------------------------------------------------
// thresholding(None) keeps the entries of sup = nms(x) that are not below max(sup) (utils/mask2samples.py:19-21).  Every
// entry of sup is x[j] (at window maxima), 0, or NaN (0 * +-inf, or a NaN sample); the global maximum of x is a window maximum.
// Hence, for a row without NaN / -inf whose maximum is > 0: max(sup) = max(x) and the kept set = { j : x[j] == max(x) }
// (+inf included: inf == inf is a window maximum, its neighbours give 0 * finite = 0).  Anything else -> literal route.
constexpr int PK_FAST_TIES = 2;                // tie positions remembered per thread
constexpr int PK_FAST_MAX = 256;               // ... and per row

struct PkFastState { float mx; int ties, t0, t1, bad; };

__device__ __forceinline__ void pk_fast_visit(PkFastState &s, float v, int p) {
    s.bad |= (v != v) | (v == -INFINITY);
    if (v > s.mx) { s.mx = v; s.ties = 0; }
    if (v == s.mx) {
        if (s.ties == 0) s.t0 = p; else if (s.ties == 1) s.t1 = p;
        ++s.ties;
    }
}

// returns the number of kept positions (written in ascending order to s_out[0 .. PK_FAST_MAX)), or -1 when the row needs
// the literal route.  Block-uniform result; contains barriers.
template <bool ALIGNED>
__device__ __forceinline__ int pk_rowmax_fast(const float *__restrict__ row, int M, float *s_red, int *s_cand, int *s_out) {
    PkFastState s;
    s.mx = -INFINITY; s.ties = 0; s.t0 = 0; s.t1 = 0; s.bad = 0;
    if (ALIGNED) {
        const float4 *r4 = reinterpret_cast<const float4 *>(row);
        const int Q = M >> 2;
        int q = threadIdx.x;
        for (; q + 3 * PK_THREADS < Q; q += 4 * PK_THREADS) {               // four 16-byte loads in flight per thread
            float4 v[4];
#pragma unroll
            for (int k = 0; k < 4; ++k) v[k] = __ldg(r4 + q + k * PK_THREADS);
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const int p = 4 * (q + k * PK_THREADS);
                pk_fast_visit(s, v[k].x, p); pk_fast_visit(s, v[k].y, p + 1); pk_fast_visit(s, v[k].z, p + 2); pk_fast_visit(s, v[k].w, p + 3);
            }
        }
        for (; q < Q; q += PK_THREADS) {
            const float4 v = __ldg(r4 + q);
            pk_fast_visit(s, v.x, 4 * q); pk_fast_visit(s, v.y, 4 * q + 1); pk_fast_visit(s, v.z, 4 * q + 2); pk_fast_visit(s, v.w, 4 * q + 3);
        }
    } else {
        for (int i = threadIdx.x; i < M; i += PK_THREADS) pk_fast_visit(s, __ldg(row + i), i);
    }
    // block maximum, "bad" flag
    float mx = s.mx;
    int bad = s.bad;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
        bad |= __shfl_xor_sync(0xffffffffu, bad, o);
    }
    const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
    if (threadIdx.x == 0) s_cand[0] = 0;
    __syncthreads();
    if (l == 0) { s_red[w] = mx; s_red[8 + w] = __int_as_float(bad); }
    __syncthreads();
    mx = s_red[0]; bad = __float_as_int(s_red[8]);
#pragma unroll
    for (int i = 1; i < PK_THREADS / 32; ++i) { mx = fmaxf(mx, s_red[i]); bad |= __float_as_int(s_red[8 + i]); }
    // candidates of the threads that saw the maximum
    if (s.mx == mx && s.ties > 0) {
        if (s.ties > PK_FAST_TIES) atomicOr(&s_cand[0], 1 << 30);
        else {
            const int at = atomicAdd(&s_cand[0], s.ties) & ((1 << 30) - 1);
            if (at + s.ties <= PK_FAST_MAX) {
                s_cand[1 + at] = s.t0;
                if (s.ties > 1) s_cand[2 + at] = s.t1;
            }
        }
    }
    __syncthreads();
    const int tot = s_cand[0];
    if (bad || !(mx > 0.0f) || tot > PK_FAST_MAX) return -1;            // also covers the tie-overflow bit
    for (int j = threadIdx.x; j < tot; j += PK_THREADS) {
        const int p = s_cand[1 + j];
        int rank = 0;
        for (int i = 0; i < tot; ++i) rank += (s_cand[1 + i] < p) ? 1 : 0;
        s_out[rank] = p;
    }
    __syncthreads();
    return tot;
}

template <bool WIN21, bool ALIGNED>
__global__ void __launch_bounds__(PK_THREADS)
peaks_kernel(const float *__restrict__ scores, int N, int M, int h, int th_is_none, float th,
             float inv_src_r, float *__restrict__ coords, int32_t *__restrict__ pos_out, int cap,
             int32_t *__restrict__ counts, unsigned long long *__restrict__ stats,
             float *__restrict__ nms_out) {
    extern __shared__ float4 pk_sm4[];
    __shared__ float s_red[16];
    __shared__ int s_warp[PK_THREADS / 32];
    const int n = blockIdx.x;
    const float *row = scores + (size_t)n * M;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const float r = inv_src_r;   // upsample_factor as float (division below is IEEE, like torch's)
    const int halo = WIN21 ? 4 * PK_HALO4 : h;
    float *sm = reinterpret_cast<float *>(pk_sm4);

    // ---- pass 1 (row-max mode only): max over the SUPPRESSED scores (thresholding receives the
    //      nms output, utils/mask2samples.py:30-31), with torch.max's NaN propagation
    float rowmax = 0.0f;
    int rownan = 0;
    const int passes = th_is_none ? 2 : 1;
    int base = 0;
    __shared__ int s_any[3];                              // "somebody keeps something" per chunk, three slots in rotation
    if (threadIdx.x < 3) s_any[threadIdx.x] = 0;          // ordered before its first use by the first chunk's barrier
    unsigned pflags = 0;
    int pp0 = 0, ci = 0;
    bool pending = false;
    // ordered per-row compaction of one chunk's kept samples (block-uniform call: contains a barrier)
    auto compact = [&](unsigned flags, int p0) {
        const int cnt = __popc(flags);
        int inc = cnt;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int t = __shfl_up_sync(0xffffffffu, inc, o);
            if (lane >= o) inc += t;
        }
        if (lane == 31) s_warp[warp] = inc;
        __syncthreads();
        int off = base + inc - cnt, tot = 0;
#pragma unroll
        for (int w = 0; w < PK_THREADS / 32; ++w) {
            const int t = s_warp[w];
            if (w < warp) off += t;
            tot += t;
        }
        while (flags) {
            const int j = __ffs(flags) - 1;
            flags &= flags - 1;
            if (off < cap) {
                if (coords) coords[(size_t)n * cap + off] = __fdiv_rn((float)(p0 + j), r);  // (:102,112)
                if (pos_out) pos_out[(size_t)n * cap + off] = p0 + j;
            }
            ++off;
        }
        base += tot;
    };
    // after the first barrier of a chunk iteration: the previous chunk's flags are visible to everybody
    auto deferred = [&]() {
        if (pending) {
            if (s_any[(ci - 1) % 3]) compact(pflags, pp0);
            pending = false;
        }
        if (threadIdx.x == 0) s_any[(ci + 1) % 3] = 0;
    };
    bool fast_done = false;
    if (th_is_none && nms_out == nullptr) {
        __shared__ int s_cand[1 + PK_FAST_MAX], s_fout[PK_FAST_MAX];
        const int kept = pk_rowmax_fast<ALIGNED>(row, M, s_red, s_cand, s_fout);
        if (kept >= 0) {
            for (int j = threadIdx.x; j < kept && j < cap; j += PK_THREADS) {
                const int p = s_fout[j];
                if (coords) coords[(size_t)n * cap + j] = __fdiv_rn((float)p, r);
                if (pos_out) pos_out[(size_t)n * cap + j] = p;
            }
            base = kept;
            fast_done = true;
        }
        __syncthreads();
    }
    for (int pass = 0; pass < passes && !fast_done; ++pass) {
        const bool measuring = th_is_none && pass == 0;
        float mx = -INFINITY;
        int nan = 0;
        float4 pre[PK_PRE];
        float4 *buf = pk_sm4;                            // WIN21 && ALIGNED: two chunk buffers, the next one filling by cp.async
        if (WIN21 && ALIGNED) {
            // ring of PK_NBUF chunk buffers, PK_NBUF - 1 chunks in flight (one cp.async group per chunk, empty past the row end)
#pragma unroll
            for (int k = 0; k < PK_NBUF - 1; ++k) {
                if (k * PK_CHUNK < M) pk_fill_async(pk_sm4 + k * PK_SMEM4, row, M, k * PK_CHUNK);
                else asm volatile("cp.async.commit_group;" ::: "memory");
            }
        } else if (WIN21) pk_load(pre, row, M, 0);
        ci = 0;
        for (int cs = 0; cs < M; cs += PK_CHUNK, ++ci) {
            if (WIN21 && ALIGNED) {
                asm volatile("cp.async.wait_group %0;" ::"n"(PK_NBUF - 2) : "memory");
                __syncthreads();                         // chunk cs visible to all; everyone is done with the previous chunk's buffer
                deferred();
                buf = pk_sm4 + (ci % PK_NBUF) * PK_SMEM4;
                if (cs + (PK_NBUF - 1) * PK_CHUNK < M)
                    pk_fill_async(pk_sm4 + ((ci + PK_NBUF - 1) % PK_NBUF) * PK_SMEM4, row, M, cs + (PK_NBUF - 1) * PK_CHUNK);
                else asm volatile("cp.async.commit_group;" ::: "memory");
            } else {
            __syncthreads();
            deferred();
            if (WIN21) {
                pk_store(pk_sm4, pre);
                if (cs + PK_CHUNK < M) pk_load(pre, row, M, cs + PK_CHUNK);      // prefetch behind this chunk's work
            } else {
                for (int i = threadIdx.x; i < PK_CHUNK + 2 * h; i += PK_THREADS) {
                    const int g = cs - h + i;
                    sm[i] = (g >= 0 && g < M) ? __ldg(row + g) : -INFINITY;
                }
            }
            __syncthreads();
            }
            float x[PK_E], sup[PK_E];
            bool skip = false;                           // warp-uniform (extraction passes of the WIN21 path only)
            const int p0 = cs + PK_E * threadIdx.x;
            if (WIN21) {
                // Pre-filter (extraction passes only): a sample can survive only if !(x < limit) - limit = th, or the row max of
                // the suppressed scores in the row-max mode - or if x = -inf (0 * -inf = NaN counts as non-zero, like the
                // reference); NaN fails `x < limit` and is covered.  A warp none of whose 512 samples qualifies skips the
                // window-max arithmetic, which otherwise (not the 4 M bytes per row) paces this kernel on mask-like data.
                bool need = (nms_out != nullptr) || measuring || (th_is_none && rownan);
                if (!need) {
                    // one NaN-propagating maximum and one minimum over the thread's 16 samples (3-input FMNMX3): the largest
                    // sample decides `!(x < limit)` for all of them (a NaN anywhere makes the maximum NaN), the smallest `-inf`
                    const float limit = th_is_none ? rowmax : th;
                    const float4 t0 = buf[pk_phys(4 * threadIdx.x + 3)], t1 = buf[pk_phys(4 * threadIdx.x + 4)];
                    const float4 t2 = buf[pk_phys(4 * threadIdx.x + 5)], t3 = buf[pk_phys(4 * threadIdx.x + 6)];
                    const float hi = pk_max3nan(pk_max3nan(pk_max3nan(t0.x, t0.y, t0.z), pk_max3nan(t0.w, t1.x, t1.y), pk_max3nan(t1.z, t1.w, t2.x)),
                                                pk_max3nan(pk_max3nan(t2.y, t2.z, t2.w), pk_max3nan(t3.x, t3.y, t3.z), t3.w), t3.w);
                    const float lo = pk_min3(pk_min3(pk_min3(t0.x, t0.y, t0.z), pk_min3(t0.w, t1.x, t1.y), pk_min3(t1.z, t1.w, t2.x)),
                                             pk_min3(pk_min3(t2.y, t2.z, t2.w), pk_min3(t3.x, t3.y, t3.z), t3.w), t3.w);
                    need = !(hi < limit) || lo == -INFINITY;
                }
                skip = !__any_sync(0xffffffffu, need);   // nothing in this warp's 512 samples can be kept: flags = 0 below
                if (!skip) pk_nms21(buf, x, sup);
            } else {
#pragma unroll 1
                for (int j = 0; j < PK_E; ++j) sup[j] = pk_nms_generic(sm, halo + PK_E * threadIdx.x + j, h);
            }
            if (nms_out != nullptr) {          // stof_nms_1d: emit the suppressed scores and nothing else
#pragma unroll
                for (int j = 0; j < PK_E; ++j)
                    if (p0 + j < M) nms_out[(size_t)n * M + p0 + j] = sup[j];
                continue;
            }
            if (measuring) {
#pragma unroll
                for (int j = 0; j < PK_E; ++j)
                    if (p0 + j < M) {
                        nan |= (sup[j] != sup[j]);
                        mx = fmaxf(mx, sup[j]);
                    }
                continue;
            }
            unsigned flags = 0;
            if (!skip) {
#pragma unroll
            for (int j = 0; j < PK_E; ++j) {
                const float s = sup[j];
                bool keep = (s != 0.0f) && (p0 + j < M);       // torch.nonzero; NaN != 0
                if (th_is_none) keep = keep && (rownan || !(s < rowmax));   // masked_fill(s < max, 0) (:21)
                else            keep = keep && !(s < th);                    // scores[scores < th] = 0 (:17)
                flags |= (keep ? 1u : 0u) << j;
            }
            }
            // The block-wide agreement "does anybody keep anything in this chunk" costs no barrier of its own: warps that do
            // raise a flag in shared memory, and the (rare) compaction runs after the NEXT chunk's first barrier.
            if (__any_sync(0xffffffffu, flags != 0) && lane == 0) s_any[ci % 3] = 1;
            pflags = flags;
            pp0 = p0;
            pending = true;
        }
        __syncthreads();
        if (pending) {
            if (s_any[(ci - 1) % 3]) compact(pflags, pp0);
            pending = false;
        }
        if (measuring) {
            const RowMax rm = block_rowmax(mx, nan, s_red);
            rowmax = rm.mx;
            rownan = rm.nan;
        }
    }
    if (nms_out != nullptr) return;
    // zero / -1 padding to the right (coords = zeros((N, K_max)), :99)
    for (int i = min(base, cap) + threadIdx.x; i < cap; i += PK_THREADS) {
        if (coords) coords[(size_t)n * cap + i] = 0.0f;
        if (pos_out) pos_out[(size_t)n * cap + i] = -1;
    }
    if (threadIdx.x == 0) {
        counts[n] = base;
        if (base) {
            atomicMax(&stats[0], (unsigned long long)base);
            atomicAdd(&stats[1], (unsigned long long)base);
        }
    }
}

// exclusive scan of counts (clamped to cap) -> offsets; single block, N up to millions is fine
__global__ void __launch_bounds__(1024)
scan_counts_kernel(const int32_t *__restrict__ counts, int N, int cap, int64_t *__restrict__ offsets) {
    __shared__ long long s_w[32];
    __shared__ long long s_carry;
    if (threadIdx.x == 0) s_carry = 0;
    __syncthreads();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int i0 = 0; i0 < N; i0 += 1024) {
        const int i = i0 + threadIdx.x;
        const long long c = (i < N) ? (long long)min(counts[i], cap) : 0;
        long long inc = c;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const long long t = __shfl_up_sync(0xffffffffu, inc, o);
            if (lane >= o) inc += t;
        }
        if (lane == 31) s_w[warp] = inc;
        __syncthreads();
        long long off = s_carry + inc - c;
        for (int w = 0; w < warp; ++w) off += s_w[w];
        if (i < N) offsets[i] = off;
        __syncthreads();
        if (threadIdx.x == 1023) s_carry = off + c;
        __syncthreads();
    }
}

__global__ void __launch_bounds__(256)
pack_indices_kernel(const int32_t *__restrict__ pos, const int32_t *__restrict__ counts, int N, int cap,
                    const int64_t *__restrict__ offsets, int64_t *__restrict__ idx) {
    const int n = blockIdx.x;
    const int c = min(counts[n], cap);
    const int64_t o = offsets[n];
    for (int k = threadIdx.x; k < c; k += blockDim.x) {
        idx[2 * (o + k) + 0] = n;
        idx[2 * (o + k) + 1] = pos[(size_t)n * cap + k];
    }
}

static int launch_peaks(const float *scores, int N, int M, int window, int th_is_none, float th,
                        int upsample_factor, float *coords, int32_t *pos, int cap, int32_t *counts,
                        int64_t *stats, float *nms_out, cudaStream_t st) {
    STOF_REQUIRE(scores && N > 0 && M > 0, STOF_E_INVALID, "stof_peaks: scores=%p N=%d M=%d", (const void *)scores, N, M);
    STOF_REQUIRE(window >= 0, STOF_E_INVALID, "stof_peaks: window=%d", window);
    const int win = window / 2 * 2 + 1;       // odd integer conversion (utils/mask2samples.py:7)
    const int h = (win - 1) / 2;
    STOF_REQUIRE(h <= PK_GEN_MAXH, STOF_E_INVALID, "stof_peaks: window %d too large (max %d)", window, 2 * PK_GEN_MAXH + 1);
    const bool aligned = (M % 4 == 0) && ((reinterpret_cast<uintptr_t>(scores) & 15) == 0);
    const float rf = (float)upsample_factor;
    unsigned long long *st64 = reinterpret_cast<unsigned long long *>(stats);
    if (h == 10) {
        const size_t smem = sizeof(float4) * PK_SMEM4 * (aligned ? PK_NBUF : 1);     // aligned rows: ring of chunk buffers (cp.async)
        static SmemAttrOnce attr;
        if (aligned) { const int rc = ensure_dyn_smem(peaks_kernel<true, true>, smem, attr); if (rc) return rc; }
        if (aligned)
            peaks_kernel<true, true><<<N, PK_THREADS, smem, st>>>(scores, N, M, h, th_is_none, th, rf, coords, pos, cap, counts, st64, nms_out);
        else
            peaks_kernel<true, false><<<N, PK_THREADS, smem, st>>>(scores, N, M, h, th_is_none, th, rf, coords, pos, cap, counts, st64, nms_out);
    } else {
        const size_t smem = sizeof(float) * (PK_CHUNK + 2 * h + 4);
        peaks_kernel<false, false><<<N, PK_THREADS, smem, st>>>(scores, N, M, h, th_is_none, th, rf, coords, pos, cap, counts, st64, nms_out);
    }
    STOF_LAUNCH_CHECK();
    return STOF_OK;
}

}  // namespace stof

extern "C" int stof_peaks(const float *scores_dev, int N, int M, int window, int th_is_none, float th,
                          int upsample_factor, float *coords_dev, int32_t *pos_dev, int cap,
                          int32_t *counts_dev, int64_t *stats_dev, void *stream) {
    STOF_REQUIRE(counts_dev && stats_dev, STOF_E_INVALID, "stof_peaks: counts/stats must not be NULL");
    STOF_REQUIRE(cap >= 0 && upsample_factor > 0, STOF_E_INVALID, "stof_peaks: cap=%d upsample_factor=%d", cap, upsample_factor);
    STOF_REQUIRE(cap == 0 || coords_dev || pos_dev, STOF_E_INVALID, "stof_peaks: cap>0 needs coords or pos");
    return stof::launch_peaks(scores_dev, N, M, window, th_is_none, th, upsample_factor, coords_dev, pos_dev, cap,
                              counts_dev, stats_dev, nullptr, (cudaStream_t)stream);
}

extern "C" int stof_nms_1d(const float *scores_dev, int N, int M, int window, float *out_dev, void *stream) {
    STOF_REQUIRE(out_dev, STOF_E_INVALID, "stof_nms_1d: out must not be NULL");
    return stof::launch_peaks(scores_dev, N, M, window, 0, 0.0f, 1, nullptr, nullptr, 0, nullptr, nullptr, out_dev,
                              (cudaStream_t)stream);
}

extern "C" int stof_peaks_indices(const int32_t *pos_dev, const int32_t *counts_dev, int N, int cap,
                                  int64_t *offsets_dev, int64_t *idx_dev, void *stream) {
    STOF_REQUIRE(pos_dev && counts_dev && offsets_dev && idx_dev && N > 0 && cap > 0, STOF_E_INVALID,
                 "stof_peaks_indices: bad arguments");
    cudaStream_t st = (cudaStream_t)stream;
    stof::scan_counts_kernel<<<1, 1024, 0, st>>>(counts_dev, N, cap, offsets_dev);
    STOF_LAUNCH_CHECK();
    stof::pack_indices_kernel<<<N, 256, 0, st>>>(pos_dev, counts_dev, N, cap, offsets_dev, idx_dev);
    STOF_LAUNCH_CHECK();
    return STOF_OK;
}
