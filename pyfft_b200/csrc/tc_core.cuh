// tcgen05 / TMEM / mbarrier building blocks for sm_100a (inline PTX, no CUTLASS).
//
// Operand convention used throughout: K-major fp16 tiles in shared memory in the canonical
// SWIZZLE_128B layout -- a tile of ROWS x 64 halves ("K atom", 128 bytes per row) is stored
// as 8-row groups of 1024 bytes; inside a group row r sits at r*128 and its eight 16-byte
// chunks are XOR-permuted with (r & 7).  K > 64 uses several atoms one after the other.
// The tile base must be 1024-byte aligned.  One tcgen05.mma consumes K = 16 halves = 32
// bytes of every row: the k-step s inside an atom is addressed by adding 32*s bytes to the
// descriptor's start address.
#pragma once
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>

namespace pyfft {
namespace tc {

// byte offset of element (row r, column k) of a K-major SW128 tile with `rows` rows (multiple of 8)
__host__ __device__ __forceinline__ uint32_t sw128_offset(int rows, int r, int k) {
    const int atom = k >> 6, kk = k & 63;
    return (uint32_t)(atom * rows * 128 + (r >> 3) * 1024 + (r & 7) * 128 + ((((kk >> 3) ^ (r & 7)) & 7) << 4) +
                      (kk & 7) * 2);
}

// shared-memory matrix descriptor: K-major, SWIZZLE_128B, 8-row group stride 1024 B
__device__ __forceinline__ uint64_t sw128_desc(uint32_t smem_addr) {
    uint64_t d = 0;
    d |= (uint64_t)((smem_addr & 0x3FFFFu) >> 4);   // start address            bits [0,14)
    d |= (uint64_t)1 << 16;                         // leading byte offset (unused for swizzled K-major)
    d |= (uint64_t)(1024 >> 4) << 32;               // stride byte offset       bits [32,46)
    d |= (uint64_t)1 << 46;                         // descriptor version 1 (sm_100)
    d |= (uint64_t)2 << 61;                         // layout type SWIZZLE_128B bits [61,64)
    return d;
}

// the same descriptor from its two halves: the high word is a constant, the low word is linear in
// the start address, so a k-step or a slot change is one 32-bit add
constexpr uint32_t SW128_DESC_HI = (uint32_t)(1024 >> 4) | (1u << 14) | (2u << 29);
__device__ __forceinline__ uint32_t sw128_desc_lo(uint32_t smem_addr) { return ((smem_addr & 0x3FFFFu) >> 4) | (1u << 16); }

// instruction descriptor for kind::f16: fp16 A and B (both K-major), fp32 accumulate
__host__ __device__ constexpr uint32_t idesc_f16(int M, int N) {
    return (1u << 4) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}

__device__ __forceinline__ void mma_f16(uint32_t d_tmem, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(d_tmem), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
        : "memory");
}

// descriptors given by their low words (see sw128_desc_lo)
__device__ __forceinline__ void mma_f16_lo(uint32_t d_tmem, uint32_t a_lo, uint32_t b_lo, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t.reg .b64 da, db;\n\t"
        "mov.b64 da, {%1, %5};\n\tmov.b64 db, {%2, %5};\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], da, db, %3, p;\n\t}"
        ::"r"(d_tmem), "r"(a_lo), "r"(b_lo), "r"(idesc), "r"(accumulate), "r"(SW128_DESC_HI)
        : "memory");
}

// A operand in tensor memory (lane = row, one 32-bit column = two consecutive K halves, 8 columns per
// K = 16 step), B from shared memory by the low word of its descriptor
__device__ __forceinline__ void mma_f16_ts_lo(uint32_t d_tmem, uint32_t a_tmem, uint32_t b_lo, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t.reg .b64 db;\n\t"
        "mov.b64 db, {%2, %5};\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], db, %3, p;\n\t}"
        ::"r"(d_tmem), "r"(a_tmem), "r"(b_lo), "r"(idesc), "r"(accumulate), "r"(SW128_DESC_HI)
        : "memory");
}

// arrive on an mbarrier when all previously issued MMAs of this thread have completed
__device__ __forceinline__ void mma_commit(uint32_t mbar_smem_addr) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(mbar_smem_addr)
                 : "memory");
}

__device__ __forceinline__ void mbar_init(uint32_t mbar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(mbar), "r"(count) : "memory");
}
__device__ __forceinline__ void fence_mbar_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }

__device__ __forceinline__ bool mbar_try_wait(uint32_t mbar, uint32_t parity) {
    uint32_t ok;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2, %3;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok)
        : "r"(mbar), "r"(parity), "r"(2000u)      // suspend-time hint (ns): sleep in hardware, do not spin
        : "memory");
    return ok != 0;
}

// try_wait without a suspend-time hint: the hardware's own (short) time limit applies.  With a hint
// the compiler emits NANOSLEEP of that length between polls, which quantises every hand-shake of a
// fine-grained pipeline.
__device__ __forceinline__ bool mbar_try_wait_nohint(uint32_t mbar, uint32_t parity) {
    uint32_t ok;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok)
        : "r"(mbar), "r"(parity)
        : "memory");
    return ok != 0;
}

__device__ __forceinline__ void mbar_arrive(uint32_t mbar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(mbar) : "memory");
}

// Wait inside a multi-role pipeline: gives up when the time-out is hit or when another role has
// raised the CTA's abort flag, so that a broken hand-shake ends the kernel instead of hanging it.
// SLEEP_NS > 0: back off between polls (roles with slack must not steal issue slots from the others).
template <int SLEEP_NS = 0>
__device__ __forceinline__ bool mbar_wait_abortable(uint32_t mbar, uint32_t parity, volatile int* abort_flag) {
#pragma unroll 1
    for (uint32_t i = 0; i < (1u << 22); ++i) {
        if (mbar_try_wait_nohint(mbar, parity)) return true;
        if (*abort_flag) return false;
        if (SLEEP_NS > 0) __nanosleep(SLEEP_NS);
    }
    *abort_flag = 1;
    return false;
}

// The same for a wait that sits on the critical ring of the pipeline (the MMA-issuing threads and
// the epilogue's wait for the stage-2 commit): the abort flag is looked at once per 32 polls, so a
// poll is try_wait + branch and the thread reacts to the phase flip as fast as it can be issued.
__device__ __forceinline__ bool mbar_wait_hot(uint32_t mbar, uint32_t parity, volatile int* abort_flag) {
#pragma unroll 1
    for (uint32_t i = 0; i < (1u << 17); ++i) {
#pragma unroll 1
        for (int j = 0; j < 32; ++j)
            if (mbar_try_wait_nohint(mbar, parity)) return true;
        if (*abort_flag) return false;
    }
    *abort_flag = 1;
    return false;
}

// Warp-collective form: lane 0 polls, the result is broadcast (the shuffle also orders the lanes'
// memory accesses after the acquire).  One poller per warp keeps the barrier unit responsive for
// the thread that issues the MMAs -- with 32 pollers per warp its own waits took ~300 cycles each.
__device__ __forceinline__ bool mbar_wait_warp(uint32_t mbar, uint32_t parity, volatile int* abort_flag) {
    int ok = 1;
    if ((threadIdx.x & 31) == 0) ok = mbar_wait_abortable(mbar, parity, abort_flag) ? 1 : 0;
    return __shfl_sync(0xffffffffu, ok, 0) != 0;
}
// register reallocation between the warpgroups of a warp-specialised kernel
template <int N>
__device__ __forceinline__ void setmaxnreg_inc() { asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;" ::"n"(N)); }
template <int N>
__device__ __forceinline__ void setmaxnreg_dec() { asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(N)); }

// Bounded wait: a wrong descriptor must not hang the GPU box.  Returns false on time-out.
__device__ __forceinline__ bool mbar_wait(uint32_t mbar, uint32_t parity) {
    for (uint32_t i = 0; i < (1u << 24); ++i)
        if (mbar_try_wait(mbar, parity)) return true;
    return false;
}

__device__ __forceinline__ void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

// one warp allocates `cols` TMEM columns (power of two >= 32); base address lands in *slot (shared)
__device__ __forceinline__ void tmem_alloc(uint32_t slot_smem_addr, uint32_t cols) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(slot_smem_addr), "r"(cols)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_free(uint32_t base, uint32_t cols) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(base), "r"(cols) : "memory");
}

// 32 consecutive fp32 columns of this thread's TMEM lane (lane = 32*(warp%4) + laneid)
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, float (&r)[32]) {
    uint32_t* u = reinterpret_cast<uint32_t*>(r);
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
        : "=r"(u[0]), "=r"(u[1]), "=r"(u[2]), "=r"(u[3]), "=r"(u[4]), "=r"(u[5]), "=r"(u[6]), "=r"(u[7]),
          "=r"(u[8]), "=r"(u[9]), "=r"(u[10]), "=r"(u[11]), "=r"(u[12]), "=r"(u[13]), "=r"(u[14]), "=r"(u[15]),
          "=r"(u[16]), "=r"(u[17]), "=r"(u[18]), "=r"(u[19]), "=r"(u[20]), "=r"(u[21]), "=r"(u[22]), "=r"(u[23]),
          "=r"(u[24]), "=r"(u[25]), "=r"(u[26]), "=r"(u[27]), "=r"(u[28]), "=r"(u[29]), "=r"(u[30]), "=r"(u[31])
        : "r"(taddr)
        : "memory");
}
// 16 consecutive fp32 columns of this thread's TMEM lane
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, float (&r)[16]) {
    uint32_t* u = reinterpret_cast<uint32_t*>(r);
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
        : "=r"(u[0]), "=r"(u[1]), "=r"(u[2]), "=r"(u[3]), "=r"(u[4]), "=r"(u[5]), "=r"(u[6]), "=r"(u[7]),
          "=r"(u[8]), "=r"(u[9]), "=r"(u[10]), "=r"(u[11]), "=r"(u[12]), "=r"(u[13]), "=r"(u[14]), "=r"(u[15])
        : "r"(taddr)
        : "memory");
}
// 8 consecutive fp32 columns of this thread's TMEM lane
__device__ __forceinline__ void tmem_ld8(uint32_t taddr, float (&r)[8]) {
    uint32_t* u = reinterpret_cast<uint32_t*>(r);
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
                 : "=r"(u[0]), "=r"(u[1]), "=r"(u[2]), "=r"(u[3]), "=r"(u[4]), "=r"(u[5]), "=r"(u[6]), "=r"(u[7])
                 : "r"(taddr)
                 : "memory");
}
// 4 consecutive fp32 columns of this thread's TMEM lane into r[0..3]
__device__ __forceinline__ void tmem_ld4(uint32_t taddr, float* r) {
    uint32_t* u = reinterpret_cast<uint32_t*>(r);
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x4.b32 {%0, %1, %2, %3}, [%4];"
                 : "=r"(u[0]), "=r"(u[1]), "=r"(u[2]), "=r"(u[3])
                 : "r"(taddr)
                 : "memory");
}
__device__ __forceinline__ void tmem_ld2(uint32_t taddr, float& a, float& b) {
    uint32_t x, y;
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x2.b32 {%0, %1}, [%2];" : "=r"(x), "=r"(y) : "r"(taddr) : "memory");
    a = __uint_as_float(x);
    b = __uint_as_float(y);
}
// 8 consecutive 32-bit columns of this thread's TMEM lane <- registers
__device__ __forceinline__ void tmem_st8(uint32_t taddr, uint32_t a, uint32_t b, uint32_t c, uint32_t d, uint32_t e,
                                         uint32_t f, uint32_t g, uint32_t h) {
    asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};"
                 ::"r"(taddr), "r"(a), "r"(b), "r"(c), "r"(d), "r"(e), "r"(f), "r"(g), "r"(h) : "memory");
}
__device__ __forceinline__ void tmem_st_wait() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

}  // namespace tc
}  // namespace pyfft
