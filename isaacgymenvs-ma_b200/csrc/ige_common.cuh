// ige_common.cuh -- shared host/device helpers for the post-physics kernels (sm_100a).
//
// Device math follows the reference's torch-CPU rounding op by op (the library is compiled with
// -fmad=false, so only the explicit __fmaf_rn calls below fuse); see DESIGN.md "Numerics".
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include "ige_b200.h"

// ---------------------------------------------------------------------------------------------
// host side: error reporting + launch accounting (api.cu)
// ---------------------------------------------------------------------------------------------
int ige_fail(int code, const char *fmt, ...);
void ige_count_launch();
int ige_check_launch(const char *what);
extern "C" int ige_use_tma();  // 1 unless IGE_NO_TMA=1 / ige_set_use_tma(0)
extern "C" int ige_fast_paths();  // 1 unless IGE_GENERIC=1 / ige_set_fast_paths(0)

extern "C" int ige_use_pdl();  // 1 unless IGE_PDL=0 / ige_set_use_pdl(0)

#define IGE_MAX_DEVICES 64   // per-device launch caches (occupancy, function attributes) are indexed by ordinal

#define IGE_REQUIRE_PTR(ptr)                                                       \
    do {                                                                           \
        if ((ptr) == nullptr) return ige_fail(IGE_ENULL, "%s is NULL", #ptr);      \
        if ((reinterpret_cast<uintptr_t>(ptr) & 15u) != 0)                         \
            return ige_fail(IGE_EALIGN, "%s is not 16-byte aligned", #ptr);        \
    } while (0)
#define IGE_OPTIONAL_PTR(ptr)                                                      \
    do {                                                                           \
        if ((ptr) != nullptr && (reinterpret_cast<uintptr_t>(ptr) & 15u) != 0)     \
            return ige_fail(IGE_EALIGN, "%s is not 16-byte aligned", #ptr);        \
    } while (0)

static inline int ige_check_rand(const IgeResetRand &r) {
    if (r.mode == IGE_RAND_INJECTED) {
        if (r.reset_rank == nullptr) return ige_fail(IGE_ENULL, "rand.reset_rank is NULL (INJECTED mode)");
        // u0/u1 may legitimately be NULL when no env resets (k == 0); the kernel never reads them then.
    } else if (r.mode != IGE_RAND_PHILOX) {
        return ige_fail(IGE_EPARAM, "rand.mode must be IGE_RAND_INJECTED or IGE_RAND_PHILOX");
    }
    return IGE_OK;
}

static inline int ige_check_list(const IgeResetList &l) {
    if (l.scratch == nullptr) return IGE_OK;
    if (l.ids == nullptr || l.count == nullptr) return ige_fail(IGE_ENULL, "list.ids / list.count is NULL");
    if ((reinterpret_cast<uintptr_t>(l.scratch) & 15u) != 0) return ige_fail(IGE_EALIGN, "list.scratch is not 16-byte aligned");
    return IGE_OK;
}

#ifdef __CUDACC__
// ---------------------------------------------------------------------------------------------
// Programmatic dependent launch (PDL).  A step is ONE short kernel (5-15 us at BASELINE.json's sizes) and steps
// follow each other on a stream or in a CUDA graph: fully serialised, every kernel pays the grid launch latency
// and its own prologue (~1 us) on top of its work.  Every kernel of the library therefore
//   * executes griddepcontrol.launch_dependents (at the top or after its main loop, see pdl_prologue below): the next
//     kernel's CTAs may be scheduled as soon as all of this kernel's CTAs have done so; they then sit in
//     griddepcontrol.wait and issue nothing, and
//   * executes griddepcontrol.wait BEFORE its first global-memory access: it returns once the previous kernel has
//     completed and its writes are visible, so the data dependence between consecutive steps is fully kept; what
//     overlaps is launch latency, argument fetch, barrier / shared-memory set-up;
// and is launched through ige_launch(), which sets cudaLaunchAttributeProgrammaticStreamSerialization (captured into
// CUDA graphs as a programmatic edge).  A kernel launched without the attribute executes both instructions as
// no-ops.  The in-kernel spin-wait protocols are unaffected: a dependent grid is only scheduled once EVERY CTA of
// its predecessor is resident or done, so it can never take a slot a running kernel is still waiting for.
// ---------------------------------------------------------------------------------------------
namespace ige {
__device__ __forceinline__ void pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
// Where a kernel triggers matters (measured, profiles/r2_pdl_trigger.txt): triggering at the very top lets the grids of
// the next 3-4 steps pile up resident behind the running one; their shared-memory reservations shrink the running
// kernel's L1, which the kernels that gather 28-byte pieces of 52-byte rows through L1 pay for (Franka PP 8192 x 2:
// 12.4 -> 16.5 us, Quadcopter 11.0 -> 15.2 us), while the TMA-staged locomotion kernel and Cartpole gain (Ant 16 384:
// 6.54 -> 5.98 us).  Policy: locomotion / Cartpole trigger at the top; every other kernel waits at the top and triggers
// after its main loop (explicitly before its tail, or implicitly by exiting), so at most ONE dependent grid is resident
// and only during the tail.
__device__ __forceinline__ void pdl_prologue() { pdl_wait(); }
__device__ __forceinline__ void pdl_prologue_early() { pdl_trigger(); pdl_wait(); }
}  // namespace ige
// SM count of the CURRENT device, cached per device ordinal (a process may drive several GPUs: a single cached value
// would size grids of device 1 with device 0's count).  *dev_out (optional) receives the ordinal, -1 if out of range.
static inline int ige_current_device_sms(int *dev_out = nullptr) {
    static int sms_of[IGE_MAX_DEVICES];
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= IGE_MAX_DEVICES) {
        if (dev_out) *dev_out = -1;
        int sms = 0;
        if (cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev < 0 ? 0 : dev) != cudaSuccess || sms < 1) sms = 1;
        return sms;
    }
    if (dev_out) *dev_out = dev;
    if (sms_of[dev] == 0) {
        int sms = 0;
        if (cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || sms < 1) sms = 1;
        sms_of[dev] = sms;
    }
    return sms_of[dev];
}
template <typename... KArgs, typename... Args>
static inline cudaError_t ige_launch_opt(bool pdl, void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem,
                                         cudaStream_t stream, Args... args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid;
    cfg.blockDim = block;
    cfg.dynamicSmemBytes = smem;
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = (pdl && ige_use_pdl()) ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, kernel, KArgs(args)...);
}
template <typename... KArgs, typename... Args>
static inline cudaError_t ige_launch(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t stream,
                                     Args... args) {
    return ige_launch_opt(true, kernel, grid, block, smem, stream, args...);
}

// ---------------------------------------------------------------------------------------------
// device math (reference: isaacgymenvs/utils/torch_jit_utils.py)
// ---------------------------------------------------------------------------------------------
namespace ige {

constexpr float kTwoPi = 6.28318530717958647692f;  // float(2*np.pi)
constexpr float kHalfPi = 1.57079632679489661923f;

// torch.norm(p=2, dim=-1) on CPU: sqrt of an fma chain
__device__ __forceinline__ float norm3(float x, float y, float z) {
    float s = __fmul_rn(x, x);
    s = __fmaf_rn(y, y, s);
    s = __fmaf_rn(z, z, s);
    return __fsqrt_rn(s);
}
__device__ __forceinline__ float norm2(float x, float y) {
    float s = __fmul_rn(x, x);
    s = __fmaf_rn(y, y, s);
    return __fsqrt_rn(s);
}
// torch.bmm (1x3)(3x1): unfused multiply-add chain (torch_jit_utils.py:88-89, :259-260)
__device__ __forceinline__ float dot3(const float *a, const float *b) {
    return __fadd_rn(__fadd_rn(__fmul_rn(a[0], b[0]), __fmul_rn(a[1], b[1])), __fmul_rn(a[2], b[2]));
}
// torch.remainder(Tensor, Scalar) (python %), torch_jit_utils.py:195
__device__ __forceinline__ float py_remainder(float a, float b) {
    float m = fmodf(a, b);
    if (m != 0.0f && ((b < 0.0f) != (m < 0.0f))) m += b;
    return m;
}
// torch_jit_utils.py:41-62, the 9-product form, xyzw
__device__ __forceinline__ void quat_mul(const float *a, const float *b, float *out) {
    const float x1 = a[0], y1 = a[1], z1 = a[2], w1 = a[3];
    const float x2 = b[0], y2 = b[1], z2 = b[2], w2 = b[3];
    const float ww = (z1 + x1) * (x2 + y2);
    const float yy = (w1 - y1) * (w2 + z2);
    const float zz = (w1 + y1) * (w2 - z2);
    const float xx = ww + yy + zz;
    const float qq = 0.5f * (xx + (z1 - x1) * (x2 - y2));
    out[3] = qq - ww + (z1 - y1) * (y2 - z2);
    out[0] = qq - xx + (x1 + w1) * (x2 + w2);
    out[1] = qq - yy + (w1 - x1) * (y2 + z2);
    out[2] = qq - zz + (z1 + y1) * (w2 - x2);
}
__device__ __forceinline__ void quat_conjugate(const float *a, float *out) {
    out[0] = -a[0]; out[1] = -a[1]; out[2] = -a[2]; out[3] = a[3];
}
// torch_jit_utils.py:80-90 (INVERSE=false) / :93-103 (INVERSE=true)
template <bool INVERSE>
__device__ __forceinline__ void quat_rotate(const float *q, const float *v, float *out) {
    const float qw = q[3];
    const float s = 2.0f * (qw * qw) - 1.0f;
    const float cr0 = q[1] * v[2] - q[2] * v[1];
    const float cr1 = q[2] * v[0] - q[0] * v[2];
    const float cr2 = q[0] * v[1] - q[1] * v[0];
    const float d = dot3(q, v);
    const float cr[3] = {cr0, cr1, cr2};
#pragma unroll
    for (int i = 0; i < 3; ++i) {
        const float a = v[i] * s;
        const float b = cr[i] * qw * 2.0f;
        const float c = q[i] * d * 2.0f;
        out[i] = INVERSE ? (a - b + c) : (a + b + c);
    }
}
// torch_jit_utils.py:65-67
__device__ __forceinline__ void normalize3(const float *x, float *out) {
    float n = norm3(x[0], x[1], x[2]);
    n = n < 1e-9f ? 1e-9f : n;
    out[0] = x[0] / n; out[1] = x[1] / n; out[2] = x[2] / n;
}
// torch_jit_utils.py:175-195; pitch is never consumed on this path
__device__ __forceinline__ void euler_roll_yaw(const float *q, float &roll, float &yaw) {
    const float qx = q[0], qy = q[1], qz = q[2], qw = q[3];
    const float sinr_cosp = 2.0f * (qw * qx + qy * qz);
    const float cosr_cosp = qw * qw - qx * qx - qy * qy + qz * qz;
    const float siny_cosp = 2.0f * (qw * qz + qx * qy);
    const float cosy_cosp = qw * qw + qx * qx - qy * qy - qz * qz;
    roll = py_remainder(atan2f(sinr_cosp, cosr_cosp), kTwoPi);
    yaw = py_remainder(atan2f(siny_cosp, cosy_cosp), kTwoPi);
}
__device__ __forceinline__ float normalize_angle(float x) { return atan2f(sinf(x), cosf(x)); }

// a / b for a loop-invariant divisor b whose reciprocal y = 1.0f / b (IEEE division) the caller hoisted.
// Two Markstein corrections (exact fma residual, then refine) make the quotient the correctly rounded
// a / b for normal-range operands -- bit for bit the `/` of the reference -- in 5 instructions instead
// of the ~15 + slow-path branch of the division sequence.  tests/test_gpu_parity.py compares the columns
// computed with it bit-exactly against the oracle's plain `/`.
__device__ __forceinline__ float div_by(float a, float b, float y) {
    float q = __fmul_rn(a, y);
    float r = __fmaf_rn(-b, q, a);
    q = __fmaf_rn(r, y, q);
    r = __fmaf_rn(-b, q, a);
    return __fmaf_rn(r, y, q);
}

// torch.sum(dim=-1) order of the CPU reference for a row of D < 32 floats held in registers:
// 8 lanes over the full groups of 8, then the tail, then the lanes (see oracle row_sum()).
template <int D, typename F>
__device__ __forceinline__ float row_sum(F elem) {
    constexpr int NV = D / 8;
    static_assert(NV < 4, "row_sum: D >= 32 needs the 4-accumulator variant");
    float lanes[8];
#pragma unroll
    for (int l = 0; l < 8; ++l) lanes[l] = 0.0f;
#pragma unroll
    for (int i = 0; i < NV; ++i)
#pragma unroll
        for (int l = 0; l < 8; ++l) lanes[l] += elem(i * 8 + l);
    float acc = 0.0f;
#pragma unroll
    for (int k = NV * 8; k < D; ++k) acc += elem(k);
    if (NV > 0) {
#pragma unroll
        for (int l = 0; l < 8; ++l) acc += lanes[l];
    }
    return acc;
}

// ---------------------------------------------------------------------------------------------
// Philox4x32-10 counter-based RNG (production reset noise; the oracle has the same function)
// ---------------------------------------------------------------------------------------------
// (the rounds stay a LOOP: unrolled, every inlined call was 1.7 KB of code that only reset rows execute, and the
// step kernels are instruction-fetch bound at small env counts -- 32 KB L1.5 instruction cache per SM)
__device__ __forceinline__ void philox4x32_10(uint32_t c[4], uint32_t k0, uint32_t k1) {
#pragma unroll 1
    for (int r = 0; r < 10; ++r) {
        const uint32_t hi0 = __umulhi(0xD2511F53u, c[0]), lo0 = 0xD2511F53u * c[0];
        const uint32_t hi1 = __umulhi(0xCD9E8D57u, c[2]), lo1 = 0xCD9E8D57u * c[2];
        const uint32_t n0 = hi1 ^ c[1] ^ k0, n2 = hi0 ^ c[3] ^ k1;
        c[0] = n0; c[1] = lo1; c[2] = n2; c[3] = lo0;
        k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
    }
}
__device__ __forceinline__ float philox_uniform(const IgeResetRand &r, uint64_t step, uint32_t row,
                                                uint32_t draw, uint32_t idx) {
    uint32_t c[4] = {row, (draw << 16) | (idx >> 2), (uint32_t)step, (uint32_t)(step >> 32)};
    philox4x32_10(c, (uint32_t)r.seed, (uint32_t)(r.seed >> 32));
    return (float)(c[idx & 3] >> 8) * (1.0f / 16777216.0f);
}
// Philox step word at kernel entry (plain load: the CTA that retires last advances it)
__device__ __forceinline__ uint64_t load_step(const IgeResetRand &r) {
    return (r.mode == IGE_RAND_PHILOX && r.step) ? *reinterpret_cast<const volatile uint64_t *>(r.step) : 0ull;
}
// uniform `j` of torch.rand call `which` (0/1) for reset row `row` with `d` columns
__device__ __forceinline__ float reset_uniform(const IgeResetRand &r, uint64_t step, int which, int row,
                                               int d, int j) {
    if (r.mode == IGE_RAND_PHILOX) return philox_uniform(r, step, (uint32_t)row, (uint32_t)which, (uint32_t)j);
    const float *u = which ? r.u1 : r.u0;
    return __ldg(u + (size_t)__ldg(r.reset_rank + row) * d + j);
}

// ---------------------------------------------------------------------------------------------
// TMA 1-D bulk copies (cp.async.bulk) + mbarrier; SASS: UBLKCP / SYNCS
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void *p) {
    return (uint32_t)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ void mbar_init(uint64_t *bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void fence_mbar_init() {
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t *bar) {   // release.cta: earlier shared-memory stores are visible to waiters
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t *bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t phase) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_%=:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra DONE_%=;\n"
        "bra WAIT_%=;\n"
        "DONE_%=:\n"
        "}\n" ::"r"(smem_u32(bar)),
        "r"(phase)
        : "memory");
}
// global -> shared, completes `bytes` on the mbarrier.  dst/src 16-B aligned, bytes % 16 == 0.
__device__ __forceinline__ void bulk_g2s(void *dst_smem, const void *src_gmem, uint32_t bytes, uint64_t *bar) {
    asm volatile(
        "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
            smem_u32(dst_smem)),
        "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar))
        : "memory");
}
// shared -> global (bulk async-group)
__device__ __forceinline__ void bulk_s2g(void *dst_gmem, const void *src_smem, uint32_t bytes) {
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst_gmem),
                 "r"(smem_u32(src_smem)), "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_read_all() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }
// make generic-proxy smem writes visible to the async proxy (before bulk_s2g)
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

// ---------------------------------------------------------------------------------------------
// episode statistics: block reduce -> one atomic per slot per CTA
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
// One warp holds NSTAT per-lane partials: reduce them and add to this CTA's statistics row with a
// single multi-lane atomic instruction (one L2 transaction on one 128-B line).
template <int NSTAT>
__device__ __forceinline__ void warp_stats_add(float *stats, const float (&v)[NSTAT], const int (&slot)[NSTAT]) {
    const int lane = threadIdx.x & 31;
    float mine = 0.0f;
    int my_slot = 0;
#pragma unroll
    for (int s = 0; s < NSTAT; ++s) {
        const float t = warp_sum(v[s]);
        if (lane == s) { mine = t; my_slot = slot[s]; }
    }
    if (lane < NSTAT) atomicAdd(stats + (blockIdx.x % IGE_STATS_ROWS) * IGE_STATS_STRIDE + my_slot, mine);
}

// ---------------------------------------------------------------------------------------------
// reset index list (IgeResetList): tile masks -> ascending ids, by the CTA that retires last
// ---------------------------------------------------------------------------------------------
// Scratch layout (uint32 words, zeroed once by the caller).  Two emission schemes share it:
//   persistent kernels (blocked tiles, one CTA scans the per-CTA counts):
//     64-bit word 0 = launch epoch, word 1 = {unused, ticket counter}, then kListMaxCtas count slots
//     {epoch + 1 : count} and kListMaxCtas prefix slots {epoch + 1 : exclusive prefix};
//   one-tile-per-warp kernels (masks + last-CTA expansion): uint32 word kListCounterWord = retirement
//     counter, ballot masks from kListMaskBase on.
constexpr int kListMaxCtas = 2040;
constexpr int kListCountBase = 2;                          // 64-bit word index
constexpr int kListPrefixBase = 2 + kListMaxCtas;          // 64-bit word index
constexpr int kListCounterWord = 2 * (2 + 2 * kListMaxCtas);
constexpr int kListMaskBase = kListCounterWord + 4;        // 16-B aligned

// uint32 words of the tagged tile slots (scheme 3 below): one 64-bit slot per 32-env tile, from kListMaskBase on;
// the plain ballot masks of the one-tile-per-warp kernels follow them
__host__ __device__ inline int list_slot_words(int num_envs) { return 2 * ((((num_envs + 31) / 32 + 3) / 4) * 4); }
__host__ __device__ inline int reset_list_scratch_words(int num_envs) {
    return kListMaskBase + list_slot_words(num_envs) + (((num_envs + 31) / 32 + 3) / 4) * 4;
}
// ---- persistent kernels: one CTA scans the per-CTA counts ----------------------------------------------------
// CTA c owns a contiguous, ascending range of envs, so its reset ids form one contiguous run of the output
// list that starts at the sum of the counts of the CTAs with a lower block index.  Every CTA reading every
// predecessor's slot is O(grid^2) loads on a few hot L2 lines (measured at 512 CTAs: 1.4 us at the tail of the
// kernel, 3 us when polled early).  Instead, O(grid) traffic and no memory fence anywhere:
//   list_cta_publish()  (one full warp per CTA, EARLY -- right after the warp counted its CTA's flags): store
//     {launch epoch + 1 : count} in the CTA's count slot and take a ticket.  Returns at once; the ticket's round
//     trip overlaps whatever the warp does next.
//   list_cta_scan()  (same warp, a little later): the CTA that drew the LAST ticket opens the next launch epoch,
//     advances the Philox step word, then reads all count slots (batched; a slot whose tag is still stale -- its
//     store may trail its CTA's ticket -- is simply re-read), scans them and writes every CTA's exclusive prefix
//     {epoch + 1 : prefix} plus the total.
//     Epoch / step word: every thread of every CTA READ both before its CTA's ticket (the caller guarantees a CTA
//     barrier between those reads and list_cta_publish), and nobody reads either again in this launch (tags are
//     compared with the register copy), so the last ticket may advance them at once.
//   list_cta_prefix_wait()  (LATE -- when the ids are about to be written): poll the CTA's OWN prefix slot; by
//     then it has long been written.
// The caller's grid must not exceed the number of co-resident CTAs (every launcher caps it with the occupancy
// API): the scanner is the last CTA to ARRIVE, so nobody waits for a CTA that is not running yet, whatever order
// the hardware dispatches CTAs in.  grid <= kListMaxCtas.
#ifndef IGE_LIST_SCAN
#define IGE_LIST_SCAN 1
#endif
struct ListTag {
    uint64_t tag;      // launch epoch + 1
    unsigned ticket;   // lane 0: this CTA's ticket (in flight until first use)
};
__device__ __forceinline__ ListTag list_cta_publish(const IgeResetList &L, int cnt) {
    volatile uint64_t *sc = reinterpret_cast<volatile uint64_t *>(L.scratch);
    const uint64_t tag = sc[0] + 1;  // (the scratch starts zeroed)
    unsigned ticket = 0;
    if ((threadIdx.x & 31) == 0) {
        sc[kListCountBase + blockIdx.x] = (tag << 32) | (uint32_t)cnt;
        ticket = atomicInc(L.scratch + 2, gridDim.x - 1u);
    }
    return ListTag{tag, ticket};
}
__device__ __forceinline__ void list_cta_scan(const IgeResetList &L, const ListTag &t, uint64_t *step_word) {
    volatile uint64_t *sc = reinterpret_cast<volatile uint64_t *>(L.scratch);
    const int lane = threadIdx.x & 31, G = (int)gridDim.x;
    if (__shfl_sync(0xffffffffu, t.ticket, 0) != (unsigned)G - 1u) return;
    if (lane == 0) {   // every CTA has consumed the epoch and the step word
        sc[0] = t.tag;
        if (step_word) *step_word += 1ull;
    }
#if IGE_LIST_SCAN
    int run = 0;
    constexpr int LB = 8;   // chunks of 32 slots in flight
    for (int c0 = 0; c0 < G; c0 += 32 * LB) {
        uint64_t v[LB];
        bool ok;
        do {
            ok = true;
#pragma unroll
            for (int k = 0; k < LB; ++k) {
                const int c = c0 + 32 * k + lane;
                v[k] = c < G ? sc[kListCountBase + c] : (t.tag << 32);
                ok &= (v[k] >> 32) == t.tag;
            }
        } while (!__all_sync(0xffffffffu, ok));
#pragma unroll
        for (int k = 0; k < LB; ++k) {
            const int c = c0 + 32 * k + lane;
            const int cnt = (int)(uint32_t)v[k];
            int incl = cnt;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const int u = __shfl_up_sync(0xffffffffu, incl, o);
                if (lane >= o) incl += u;
            }
            if (c < G) sc[kListPrefixBase + c] = (t.tag << 32) | (uint32_t)(run + incl - cnt);
            run += __shfl_sync(0xffffffffu, incl, 31);
        }
    }
    if (lane == 0) *L.count = run;
#endif
}
// exclusive prefix of this CTA (number of ids emitted by the CTAs with a lower block index); full warp
__device__ __forceinline__ int list_cta_prefix_wait(const IgeResetList &L, const ListTag &t) {
    const volatile uint64_t *sc = reinterpret_cast<const volatile uint64_t *>(L.scratch);
#if !IGE_LIST_SCAN
    {   // (round-1 behaviour, kept for A/B timing) batched polls of every predecessor's count slot
        const int lane = threadIdx.x & 31;
        int sum = 0;
        constexpr int LB = 8;
        for (int c0 = lane; c0 < (int)blockIdx.x; c0 += 32 * LB) {
            uint64_t v[LB];
            bool ok;
            do {
                ok = true;
#pragma unroll
                for (int k = 0; k < LB; ++k) {
                    const int c = c0 + 32 * k;
                    v[k] = c < (int)blockIdx.x ? sc[kListCountBase + c] : (t.tag << 32);
                    ok &= (v[k] >> 32) == t.tag;
                }
            } while (!ok);
#pragma unroll
            for (int k = 0; k < LB; ++k) sum += (int)(uint32_t)v[k];
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
        if (blockIdx.x == gridDim.x - 1 && lane == 0) *L.count = sum + (int)(uint32_t)sc[kListCountBase + blockIdx.x];
        return sum;
    }
#endif
    uint64_t v;
    do { v = sc[kListPrefixBase + blockIdx.x]; } while ((v >> 32) != t.tag);
    return (int)(uint32_t)v;
}
// Ascending ids of one 32-env ballot `mask` whose first env is `base`, written from list position `off` on.
// One lane per call (the caller distributes tiles over lanes); returns the next free position.
__device__ __forceinline__ int list_emit_mask(const IgeResetList &L, int off, uint32_t mask, int base) {
    while (mask) {
        L.ids[off++] = base + __ffs(mask) - 1;
        mask &= mask - 1;
    }
    return off;
}

// ---- persistent kernels, scheme 3: tagged tile slots, one scanner CTA, no worker ever waits -------------------
// Any cross-CTA hand-off costs L2 round trips (~0.7 us each on B200).  Schemes that make every CTA wait for a prefix
// put three or four of them, chained, at the tail of a 5-8 us kernel.  Here the workers only STORE:
//   worker, per 32-env tile:  list_publish_slot(): one 64-bit store {launch epoch + 1 : ballot of the entry flags}
//     as soon as the tile's flags are in registers.  Fire and forget -- no ticket, no fence, no poll.
//   scanner = one extra CTA (the last block index):  list_scanner_cta() polls the slots in tile order (the worker
//     CTAs take tile = block + k * workers, so consecutive tiles are published at about the same time), scans the
//     popcounts and writes the ascending ids itself; after the last slot it stores the count, opens the next epoch
//     and advances the Philox step word.  The ids of a run of tiles are complete about one L2 round trip after its
//     last slot was published, i.e. while the workers are still storing that tile's outputs.
// In this basic form nobody waits for the scanner, so the grid needs no co-residency guarantee (the launchers still
// keep workers + 1 <= resident CTAs so that the scanner starts with the kernel).  EXCEPTION: the paired Franka rows
// kernel (Combine / PP, franka_ma.cu) adds a gate -- its workers DO wait for the scanner's epoch store before they
// overwrite the entry flags, and the scanner waits for every worker's slots -- so that caller needs every CTA
// resident at once; its launcher keeps one SM's worth of CTAs free for kernels of other streams (statistics
// exchange, NCCL) that may be resident beside it.  Epoch / step word: every worker CTA read
// both before its first publish (the caller guarantees a CTA barrier in between), and the scanner advances them only
// after it has seen every slot of this launch.
__device__ __forceinline__ uint32_t list_epoch_tag(const IgeResetList &L) {
    return (uint32_t)(*reinterpret_cast<const volatile uint64_t *>(L.scratch)) + 1u;
}
__device__ __forceinline__ void list_publish_slot(const IgeResetList &L, uint32_t tag, int tile, uint32_t mask) {
    asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(__cvta_generic_to_global(
                     reinterpret_cast<uint64_t *>(L.scratch + kListMaskBase) + tile)), "l"(((uint64_t)tag << 32) | mask) : "memory");
}
// entry flag of env `e` as published by its owner (cross-CTA read of ENTRY flags without touching reset_buf, which
// the owner may already have overwritten); spins until the owner's slot of this launch is there
__device__ __forceinline__ bool list_slot_bit(const IgeResetList &L, uint32_t tag, int e) {
    const volatile uint64_t *slots = reinterpret_cast<const volatile uint64_t *>(L.scratch + kListMaskBase);
    uint64_t v;
    do { v = slots[e >> 5]; } while ((uint32_t)(v >> 32) != tag);
    return ((uint32_t)v >> (e & 31)) & 1u;
}
// The scanner works in UNITS of 256 consecutive tiles: lane l owns the 8 contiguous slots [8 l, 8 l + 8) (four 16-B
// loads, all in flight), so ONE warp scan of the lane totals orders the whole unit.  Unit u belongs to warp
// u % (T / 32): T / 32 units are polled concurrently (under load an L2 round trip is 1-3 us; a scanner with one unit
// in flight falls behind the workers at 1M envs).  The running offset passes from unit to unit through one 64-bit
// shared-memory word, BEFORE the unit's ids are written.  The code is kept SMALL on purpose: the scanner shares
// its SM's instruction cache with five worker CTAs, and a long unrolled scan measured 11 us per unit in fetch
// stalls alone.
// Slot accesses are gpu-scope relaxed (volatile would be system-scope strong accesses through generic addressing).
constexpr int kListUnitSlots = 8;                       // per lane
constexpr int kListUnitTiles = 32 * kListUnitSlots;     // per unit
__device__ __forceinline__ uint64_t ld_relaxed_gpu_u64(const uint64_t *p) {
    uint64_t v;
    asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(__cvta_generic_to_global(p)) : "memory");
    return v;
}
__device__ __forceinline__ void ld_relaxed_gpu_2xu64(const uint64_t *p, uint64_t &a, uint64_t &b) {
    asm volatile("ld.relaxed.gpu.global.v2.u64 {%0, %1}, [%2];" : "=l"(a), "=l"(b) : "l"(__cvta_generic_to_global(p)) : "memory");
}
#ifdef IGE_SCAN_DEBUG
// [0][0] = min worker start, [0][1] = max worker end, [0][2] = scanner start; [1 + u][..] per unit: poll start,
// polls, all valid, chain passed, ids written, scanned, spun
__device__ unsigned long long g_scan_dbg[80][8];
__device__ __forceinline__ unsigned long long gtime() { unsigned long long t; asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t)); return t; }
#define SCAN_DBG(u, k, v) do { if ((threadIdx.x & 31) == 0 && (u) < 79) g_scan_dbg[1 + (u)][k] = (v); } while (0)
#else
#define SCAN_DBG(u, k, v) do { } while (0)
#endif
// S = slots per lane (unit = 32 S tiles).  CTAs with few warps (T = 64: the Franka MA rows kernel) take a larger S so that
// large env counts do not queue up behind two units in flight (262 144 envs x 2 agents: 195 -> see DESIGN.md us).
template <int T, int S = kListUnitSlots>
static __device__ __noinline__ int list_scanner_cta(const IgeResetList L, const int num_tiles, uint64_t *step_word) {
    constexpr int W = T / 32;
    constexpr int kUnitTiles = 32 * S;
    __shared__ volatile unsigned long long s_chain;   // {units done : ids so far}
    if (threadIdx.x == 0) s_chain = 0ull;
    __syncthreads();
#ifdef IGE_SCAN_DEBUG
    if (threadIdx.x == 0) g_scan_dbg[0][2] = gtime();
#endif
#if !defined(IGE_EXP_SCAN_NOP)
    // (the slot region is padded to a multiple of 4 tiles and zeroed once: slots past num_tiles are never valid, so
    // they are not loaded)
    const uint64_t *slots = reinterpret_cast<const uint64_t *>(L.scratch + kListMaskBase);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const uint32_t tag = (uint32_t)ld_relaxed_gpu_u64(reinterpret_cast<const uint64_t *>(L.scratch)) + 1u;
    const uint64_t done = (uint64_t)tag << 32;   // value of a slot with no reset (also stands in for slots past the end)
    const int units = (num_tiles + kUnitTiles - 1) / kUnitTiles;
#pragma unroll 1
    for (int u = warp; u < units; u += W) {
        const int lo = u * kUnitTiles + lane * S;
        uint64_t v[S];
        int polls = 0;
        (void)polls;
        SCAN_DBG(u, 0, gtime());
#pragma unroll
        for (int i = 0; i < S; ++i) v[i] = (lo + i < num_tiles) ? 0ull : done;
        for (;;) {   // (re-)load ALL stale pairs per spin: independent loads, one round trip per spin
            bool ok = true;
#pragma unroll
            for (int i = 0; i < S; i += 2)
                if ((uint32_t)(v[i] >> 32) != tag || (uint32_t)(v[i + 1] >> 32) != tag) {
                    if (lo + i + 1 < num_tiles) ld_relaxed_gpu_2xu64(slots + lo + i, v[i], v[i + 1]);
                    else v[i] = ld_relaxed_gpu_u64(slots + lo + i);   // odd tail (lo + i < num_tiles here)
                    ok = false;
                }
            if (ok) break;
            ++polls;
        }
        __syncwarp();
        SCAN_DBG(u, 1, (unsigned long long)polls);
        SCAN_DBG(u, 2, gtime());
        int cnt = 0;
#pragma unroll
        for (int i = 0; i < S; ++i) cnt += __popc((uint32_t)v[i]);
        int incl = cnt;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int t = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += t;
        }
        int run = 0;
        SCAN_DBG(u, 5, gtime());
        if (lane == 31) {   // incl == the unit's total.  Hand-off word: ONE 64-bit shared-memory word, so no fence
            unsigned long long w;
            do { w = s_chain; } while ((int)(w >> 32) != u);
            SCAN_DBG(u, 6, gtime());
            run = (int)(uint32_t)w;
            s_chain = ((unsigned long long)(u + 1) << 32) | (uint32_t)(run + incl);
            if (u == units - 1) {   // every slot of this launch has been seen
                *L.count = run + incl;
                *reinterpret_cast<volatile uint64_t *>(L.scratch) = tag;
                if (step_word) *step_word += 1ull;
            }
        }
        run = __shfl_sync(0xffffffffu, run, 31);
        SCAN_DBG(u, 3, gtime());
        int off = run + incl - cnt;
        if (cnt > 0) {
#pragma unroll
            for (int i = 0; i < S; ++i) {
                uint32_t m = (uint32_t)v[i];
                const int e0 = (lo + i) * 32;
#pragma unroll 1
                while (m) {   // (not unrolled: the compiler's 4x unrolling of these 8 loops was 16 KB of code)
                    L.ids[off++] = e0 + __ffs(m) - 1;
                    m &= m - 1;
                }
            }
        }
        __syncwarp();
        SCAN_DBG(u, 4, gtime());
    }
#else
    if (threadIdx.x == 0) { *L.count = 0; *reinterpret_cast<volatile uint64_t *>(L.scratch) += 1; if (step_word) *step_word += 1ull; }
#endif
    __syncthreads();
    return (int)(uint32_t)s_chain;   // number of ids (same value in every thread)
}

// One thread publishes the ballot mask of 32-env tile `tile` (bit i = env tile*32+i was reset).
// No fence here: finish_reset_list() fences once per publishing thread.
__device__ __forceinline__ void publish_tile_mask(const IgeResetList &L, int num_envs, int tile, uint32_t mask) {
    L.scratch[kListMaskBase + list_slot_words(num_envs) + tile] = mask;
}
// finish_reset_list() is called by EVERY thread of EVERY CTA once, after the CTA's masks are published
// (`i_published`: this thread wrote masks).  The CTA that arrives last expands all masks (blocked
// partition, batched vector loads, block scan) into L.ids / L.count, re-zeroes the counter and
// advances the Philox step word.
// (kept out of line: it runs in ONE CTA per launch and must not inflate the step kernels' registers)
static __device__ __noinline__ void expand_reset_list(const IgeResetList L, const int num_envs, uint64_t *step) {
    __shared__ int s_wtot[32];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nthr = blockDim.x, nwarps = (nthr + 31) >> 5;
    __threadfence();
    const int tiles4 = ((num_envs + 31) / 32 + 3) / 4;
    const uint4 *m4 = reinterpret_cast<const uint4 *>(L.scratch + kListMaskBase + list_slot_words(num_envs));
    const int per = (tiles4 + nthr - 1) / nthr;
    const int lo = min(tiles4, tid * per), hi = min(tiles4, lo + per);
    constexpr int B = 4;  // loads in flight per thread
    int cnt = 0;
    for (int i0 = lo; i0 < hi; i0 += B) {
        uint4 m[B];
#pragma unroll
        for (int k = 0; k < B; ++k) m[k] = (i0 + k < hi) ? __ldcg(m4 + i0 + k) : make_uint4(0, 0, 0, 0);
#pragma unroll
        for (int k = 0; k < B; ++k) cnt += __popc(m[k].x) + __popc(m[k].y) + __popc(m[k].z) + __popc(m[k].w);
    }
    int incl = cnt;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int t = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += t;
    }
    if (lane == 31) s_wtot[warp] = incl;
    __syncthreads();
    int woff = 0, total = 0;
    for (int w = 0; w < nwarps; ++w) {
        const int t = s_wtot[w];
        if (w < warp) woff += t;
        total += t;
    }
    int off = woff + incl - cnt;
    if (cnt > 0) {
        for (int i0 = lo; i0 < hi; i0 += B) {
            uint4 m[B];
#pragma unroll
            for (int k = 0; k < B; ++k) m[k] = (i0 + k < hi) ? __ldcg(m4 + i0 + k) : make_uint4(0, 0, 0, 0);
#pragma unroll
            for (int k = 0; k < B; ++k) {
                const uint32_t w4[4] = {m[k].x, m[k].y, m[k].z, m[k].w};
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    uint32_t mw = w4[q];
                    const int base = ((i0 + k) * 4 + q) * 32;
                    while (mw) {
                        L.ids[off++] = base + __ffs(mw) - 1;
                        mw &= mw - 1;
                    }
                }
            }
        }
    }
    if (tid == 0) {
        *L.count = total;
        L.scratch[kListCounterWord] = 0u;
        if (step) *step += 1ull;
    }
}
__device__ __forceinline__ void finish_reset_list(const IgeResetList &L, int num_envs, uint64_t *step,
                                                  bool i_published) {
    __shared__ int s_last;
    if (i_published) __threadfence();  // my masks are visible device-wide before this CTA counts as retired
    __syncthreads();
    if (threadIdx.x == 0) s_last = atomicAdd(&L.scratch[kListCounterWord], 1u) == gridDim.x - 1;
    __syncthreads();
    if (s_last) expand_reset_list(L, num_envs, step);
}

}  // namespace ige
#endif  // __CUDACC__
