// Thin inline-PTX wrappers (sm_100a) used by the fused kernels.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace bq {
namespace ptx {

__device__ __forceinline__ uint64_t globaltimer() {
    uint64_t t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}

// ---- system-scope flag / slot traffic over NVLink peer mappings ----------
__device__ __forceinline__ uint64_t ld_acquire_sys(const uint64_t* p) {
    uint64_t v;
    asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ uint64_t ld_acquire_gpu(const uint64_t* p) {
    uint64_t v;
    asm volatile("ld.acquire.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ unsigned int ld_acquire_gpu_u32(const unsigned int* p) {
    unsigned int v;
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_release_gpu(uint64_t* p, uint64_t v) {
    asm volatile("st.release.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ void red_release_sys_add(uint64_t* p, uint64_t v) {
    asm volatile("red.release.sys.global.add.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ void st_relaxed_sys_f64x2(double* p, double a, double b) {
    asm volatile("st.relaxed.sys.global.v2.f64 [%0], {%1, %2};" ::"l"(p), "d"(a), "d"(b)
                 : "memory");
}
__device__ __forceinline__ void ld_relaxed_sys_f64x2(const double* p, double& a, double& b) {
    asm volatile("ld.relaxed.sys.global.v2.f64 {%0, %1}, [%2];" : "=d"(a), "=d"(b) : "l"(p)
                 : "memory");
}
__device__ __forceinline__ void st_relaxed_sys_f64(double* p, double a) {
    asm volatile("st.relaxed.sys.global.f64 [%0], %1;" ::"l"(p), "d"(a) : "memory");
}
__device__ __forceinline__ double ld_relaxed_sys_f64(const double* p) {
    double a;
    asm volatile("ld.relaxed.sys.global.f64 %0, [%1];" : "=d"(a) : "l"(p) : "memory");
    return a;
}
// ---- LL words: 32 bits of payload | 32-bit call number, one 8-byte store each -------------
__device__ __forceinline__ void st_relaxed_sys_u64x2(uint64_t* p, uint64_t a, uint64_t b) {
    asm volatile("st.relaxed.sys.global.v2.u64 [%0], {%1, %2};" ::"l"(p), "l"(a), "l"(b) : "memory");
}
__device__ __forceinline__ uint64_t ld_relaxed_sys_u64(const uint64_t* p) {
    uint64_t v;
    asm volatile("ld.relaxed.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_relaxed_gpu_u64x2(uint64_t* p, uint64_t a, uint64_t b) {
    asm volatile("st.relaxed.gpu.global.v2.u64 [%0], {%1, %2};" ::"l"(p), "l"(a), "l"(b) : "memory");
}
__device__ __forceinline__ void ld_relaxed_gpu_u64x2(const uint64_t* p, uint64_t& a, uint64_t& b) {
    asm volatile("ld.relaxed.gpu.global.v2.u64 {%0, %1}, [%2];" : "=l"(a), "=l"(b) : "l"(p) : "memory");
}
__device__ __forceinline__ void ld_relaxed_sys_u64x2(const uint64_t* p, uint64_t& a, uint64_t& b) {
    asm volatile("ld.relaxed.sys.global.v2.u64 {%0, %1}, [%2];" : "=l"(a), "=l"(b) : "l"(p) : "memory");
}
__device__ __forceinline__ void st_release_sys_u64(uint64_t* p, uint64_t v) {
    asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
// One double as two LL words (low half | flag, high half | flag): a 16-byte store.  SYS = the
// word crosses NVLink (peer memory) or is read by a peer; otherwise it stays inside this GPU.
template <bool SYS>
__device__ __forceinline__ void ll_put(uint64_t* dst, double v, uint32_t flag) {
    const uint64_t b = uint64_t(__double_as_longlong(v)), f = uint64_t(flag) << 32;
    if (SYS)
        st_relaxed_sys_u64x2(dst, (b & 0xffffffffull) | f, (b >> 32) | f);
    else
        st_relaxed_gpu_u64x2(dst, (b & 0xffffffffull) | f, (b >> 32) | f);
}
// True (and v set) when both words carry `flag`.
template <bool SYS>
__device__ __forceinline__ bool ll_get(const uint64_t* src, uint32_t flag, double& v) {
    uint64_t a, b;
    if (SYS)
        ld_relaxed_sys_u64x2(src, a, b);
    else
        ld_relaxed_gpu_u64x2(src, a, b);
    if (uint32_t(a >> 32) != flag || uint32_t(b >> 32) != flag) return false;
    v = __longlong_as_double((long long)((a & 0xffffffffull) | (b << 32)));
    return true;
}

// NVLS: one store to a multicast address, replicated by the NVSwitch into every bound GPU's memory.
__device__ __forceinline__ void multimem_st_u64(uint64_t* mc, uint64_t v) {
    asm volatile("multimem.st.relaxed.sys.global.u64 [%0], %1;" ::"l"(mc), "l"(v) : "memory");
}
// 16 bytes (two doubles) to a multicast address: the switch writes them into every GPU's copy.
// (multimem.st has no .f64 vector form; the two doubles travel as four 32-bit lanes, bit for bit.)
__device__ __forceinline__ void multimem_st_f64x2(double* mc, double a, double b) {
    const unsigned long long ua = (unsigned long long)__double_as_longlong(a), ub = (unsigned long long)__double_as_longlong(b);
    asm volatile("multimem.st.weak.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(mc),
                 "f"(__uint_as_float(unsigned(ua))), "f"(__uint_as_float(unsigned(ua >> 32))),
                 "f"(__uint_as_float(unsigned(ub))), "f"(__uint_as_float(unsigned(ub >> 32)))
                 : "memory");
}
__device__ __forceinline__ void multimem_red_add_u64(uint64_t* mc, uint64_t v) {
    asm volatile("multimem.red.release.sys.global.add.u64 [%0], %1;" ::"l"(mc), "l"(v) : "memory");
}
// 4 LL words of a dd pair tagged with `flag` (never 0 for a live call: flag = call number + 1).
__device__ __forceinline__ void ll_pack(double hi, double lo, uint32_t flag, uint64_t (&w)[4]) {
    const uint64_t h = uint64_t(__double_as_longlong(hi)), l = uint64_t(__double_as_longlong(lo));
    const uint64_t f = uint64_t(flag) << 32;
    w[0] = (h & 0xffffffffull) | f;
    w[1] = (h >> 32) | f;
    w[2] = (l & 0xffffffffull) | f;
    w[3] = (l >> 32) | f;
}
__device__ __forceinline__ void ll_store(uint64_t* dst, double hi, double lo, uint32_t flag) {
    uint64_t w[4];
    ll_pack(hi, lo, flag, w);
    st_relaxed_sys_u64x2(dst, w[0], w[1]);
    st_relaxed_sys_u64x2(dst + 2, w[2], w[3]);
}
__device__ __forceinline__ void ll_store_multicast(uint64_t* mc_dst, double hi, double lo, uint32_t flag) {
    uint64_t w[4];
    ll_pack(hi, lo, flag, w);
#pragma unroll
    for (int i = 0; i < 4; ++i) multimem_st_u64(mc_dst + i, w[i]);
}

__device__ __forceinline__ void fence_sys() { asm volatile("fence.acq_rel.sys;" ::: "memory"); }

// 256-bit global store (sm_100+): one full 32-byte sector per thread.
__device__ __forceinline__ void st_f64x4(double* p, double a, double b, double c, double d) {
    asm volatile("st.global.v4.f64 [%0], {%1, %2, %3, %4};" ::"l"(p), "d"(a), "d"(b), "d"(c), "d"(d)
                 : "memory");
}
__device__ __forceinline__ void st_f64x2(double* p, double a, double b) {
    asm volatile("st.global.v2.f64 [%0], {%1, %2};" ::"l"(p), "d"(a), "d"(b) : "memory");
}
// Streaming variant for outputs that are written once and read by another kernel.
__device__ __forceinline__ void st_cs_f64x2(double* p, double a, double b) {
    asm volatile("st.global.cs.v2.f64 [%0], {%1, %2};" ::"l"(p), "d"(a), "d"(b) : "memory");
}

// ---- mbarrier + 1-D TMA bulk copy (global -> shared) ----------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) {
    return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)),
                 "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_%=:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra DONE_%=;\n"
        "bra WAIT_%=;\n"
        "DONE_%=:\n"
        "}\n" ::"r"(smem_u32(bar)),
        "r"(parity)
        : "memory");
}
// bytes % 16 == 0, both addresses 16-byte aligned.  SASS: UBLKCP.
__device__ __forceinline__ void tma_bulk_g2s(void* smem_dst, const void* gmem_src, uint32_t bytes,
                                             uint64_t* bar) {
    asm volatile(
        "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::
            "r"(smem_u32(smem_dst)),
        "l"(gmem_src), "r"(bytes), "r"(smem_u32(bar))
        : "memory");
}

// ---- thread-block cluster -------------------------------------------------
__device__ __forceinline__ uint32_t cluster_ctarank() {
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
    return r;
}
__device__ __forceinline__ uint32_t cluster_nctarank() {
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_nctarank;" : "=r"(r));
    return r;
}
__device__ __forceinline__ void cluster_arrive() {
    asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
}
__device__ __forceinline__ void cluster_wait() {
    asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ void cluster_sync_all() {
    asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
    asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// Address of `p` (own shared memory) as seen in CTA `rank` of the cluster.
__device__ __forceinline__ uint32_t mapa(const void* p, uint32_t rank) {
    uint32_t out;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(out) : "r"(smem_u32(p)), "r"(rank));
    return out;
}
__device__ __forceinline__ void st_dsmem_f64x2(uint32_t addr, double a, double b) {
    asm volatile("st.shared::cluster.v2.f64 [%0], {%1, %2};" ::"r"(addr), "d"(a), "d"(b) : "memory");
}

}  // namespace ptx
}  // namespace bq
