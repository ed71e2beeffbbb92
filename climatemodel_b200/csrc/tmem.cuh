// Tensor memory (TMEM, 256 KB per SM on sm_100a) used as per-lane private scratch.
//
// TMEM is 128 lanes x 512 columns of 32 bits; warp w of a CTA reaches lanes 32 (w % 4) .. + 31, and
// with the 32x32b access shape thread t of the warp reads / writes consecutive columns of lane
// 32 (w % 4) + t.  That is exactly the access pattern of the flux kernel's t-store (every chain = lane
// parks its own layer transmissivities and replays them in the opposite direction), so the t-store
// can live here instead of in shared memory: one LDTM / STTM moves a block of kU values per lane.
#pragma once
#include <cstdint>

namespace cmb {
namespace tmem {

// One warp allocates n_cols (power of two, 32..512) columns for the CTA; the base address lands
// in *slot (shared memory).  The same warp frees them with release().
__device__ __forceinline__ void acquire(uint32_t *slot, uint32_t n_cols) {
    const uint32_t dst = (uint32_t)__cvta_generic_to_shared(slot);
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(dst), "r"(n_cols) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void release(uint32_t base, uint32_t n_cols) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(base), "r"(n_cols) : "memory");
}
__device__ __forceinline__ void fence_before_sync() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void fence_after_sync() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void wait_st() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }

// Address of column c in the lane quarter of warp `warp`.
__device__ __forceinline__ uint32_t address(uint32_t base, int warp, int col) {
    return base + ((uint32_t)((warp & 3) * 32) << 16) + (uint32_t)col;
}

// Block load / store of NW consecutive columns per lane (NW = 8 or 16).  The loaded registers are
// only valid after wait_ld(), which takes them as in-out operands so that no use is scheduled earlier.
template <int NW>
__device__ __forceinline__ void ld(uint32_t addr, uint32_t (&r)[NW]) {
    static_assert(NW == 8 || NW == 16, "block of 8 or 16 columns");
    if constexpr (NW == 16) {
        asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
                     : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
                       "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
                     : "r"(addr));
    } else {
        asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
                     : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7])
                     : "r"(addr));
    }
}
template <int NW>
__device__ __forceinline__ void wait_ld(uint32_t (&r)[NW]) {
    if constexpr (NW == 16) {
        asm volatile("tcgen05.wait::ld.sync.aligned;"
                     : "+r"(r[0]), "+r"(r[1]), "+r"(r[2]), "+r"(r[3]), "+r"(r[4]), "+r"(r[5]), "+r"(r[6]), "+r"(r[7]),
                       "+r"(r[8]), "+r"(r[9]), "+r"(r[10]), "+r"(r[11]), "+r"(r[12]), "+r"(r[13]), "+r"(r[14]), "+r"(r[15])
                     :
                     : "memory");
    } else {
        asm volatile("tcgen05.wait::ld.sync.aligned;"
                     : "+r"(r[0]), "+r"(r[1]), "+r"(r[2]), "+r"(r[3]), "+r"(r[4]), "+r"(r[5]), "+r"(r[6]), "+r"(r[7])
                     :
                     : "memory");
    }
}
template <int NW>
__device__ __forceinline__ void st(uint32_t addr, const uint32_t (&r)[NW]) {
    static_assert(NW == 8 || NW == 16, "block of 8 or 16 columns");
    if constexpr (NW == 16) {
        asm volatile("tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16};"
                     :
                     : "r"(addr), "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]),
                       "r"(r[8]), "r"(r[9]), "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15])
                     : "memory");
    } else {
        asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};"
                     :
                     : "r"(addr), "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7])
                     : "memory");
    }
}

// Two or four consecutive columns per lane (one or two FP64 values): small operand groups that ptxas can
// keep in the registers the values already live in.
template <int NW>
__device__ __forceinline__ void ld_small(uint32_t addr, uint32_t (&r)[NW]) {
    static_assert(NW == 2 || NW == 4, "2 or 4 columns");
    if constexpr (NW == 4) {
        asm volatile("tcgen05.ld.sync.aligned.32x32b.x4.b32 {%0, %1, %2, %3}, [%4];"
                     : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3])
                     : "r"(addr));
    } else {
        asm volatile("tcgen05.ld.sync.aligned.32x32b.x2.b32 {%0, %1}, [%2];" : "=r"(r[0]), "=r"(r[1]) : "r"(addr));
    }
}
__device__ __forceinline__ void st_one(uint32_t addr, uint32_t r) {      // one column per lane (an FP32 value)
    asm volatile("tcgen05.st.sync.aligned.32x32b.x1.b32 [%0], {%1};" ::"r"(addr), "r"(r) : "memory");
}
template <int NW>
__device__ __forceinline__ void st_small(uint32_t addr, const uint32_t (&r)[NW]) {
    static_assert(NW == 2 || NW == 4, "2 or 4 columns");
    if constexpr (NW == 4) {
        asm volatile("tcgen05.st.sync.aligned.32x32b.x4.b32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]) : "memory");
    } else {
        asm volatile("tcgen05.st.sync.aligned.32x32b.x2.b32 [%0], {%1, %2};" ::"r"(addr), "r"(r[0]), "r"(r[1]) : "memory");
    }
}
// Wait for every outstanding tcgen05.ld of the thread; the loaded words are in-out operands so that no
// use is scheduled before the wait.
template <int NW>
__device__ __forceinline__ void wait_ld_all(uint32_t (&r)[NW]) {
    static_assert(NW == 16 || NW == 8, "a block of words");
    if constexpr (NW == 16) {
        asm volatile("tcgen05.wait::ld.sync.aligned;"
                     : "+r"(r[0]), "+r"(r[1]), "+r"(r[2]), "+r"(r[3]), "+r"(r[4]), "+r"(r[5]), "+r"(r[6]), "+r"(r[7]),
                       "+r"(r[8]), "+r"(r[9]), "+r"(r[10]), "+r"(r[11]), "+r"(r[12]), "+r"(r[13]), "+r"(r[14]), "+r"(r[15])
                     :
                     : "memory");
    } else {
        asm volatile("tcgen05.wait::ld.sync.aligned;"
                     : "+r"(r[0]), "+r"(r[1]), "+r"(r[2]), "+r"(r[3]), "+r"(r[4]), "+r"(r[5]), "+r"(r[6]), "+r"(r[7])
                     :
                     : "memory");
    }
}

// Transposing load, shape 16x256b.x2: the warp reads 16 columns of 16 consecutive TMEM lanes starting at the lane
// in addr (its quarter's first lane, or that + 16); thread t receives columns 2 (t % 4), 2 (t % 4) + 1 of lane
// t / 4 (r[0], r[1]) and of lane t / 4 + 8 (r[2], r[3]), and columns 8 + 2 (t % 4), 9 + 2 (t % 4) of the same two
// lanes (r[4..7]).  With FP64 values parked by st<16> (value u = columns 2 u, 2 u + 1 of the owner's lane) that is
// value t % 4 (r[0..3]) and value t % 4 + 4 (r[4..7]) of lanes t / 4 and t / 4 + 8.
__device__ __forceinline__ void ld_16x256b_x2(uint32_t addr, uint32_t (&r)[8]) {
    asm volatile("tcgen05.ld.sync.aligned.16x256b.x2.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7])
                 : "r"(addr));
}
// The same shape over 8 columns (.x1): columns 2 (t % 4), 2 (t % 4) + 1 of lanes t / 4 (r[0], r[1]) and t / 4 + 8
// (r[2], r[3]) -- FP32 values 2 (t % 4) and 2 (t % 4) + 1 of those lanes when 8 values were parked per lane.
__device__ __forceinline__ void ld_16x256b_x1(uint32_t addr, uint32_t (&r)[4]) {
    asm volatile("tcgen05.ld.sync.aligned.16x256b.x1.b32 {%0, %1, %2, %3}, [%4];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3])
                 : "r"(addr));
}
__device__ __forceinline__ void wait_ld_pair(uint32_t (&a)[4], uint32_t (&b)[4]) {
    asm volatile("tcgen05.wait::ld.sync.aligned;"
                 : "+r"(a[0]), "+r"(a[1]), "+r"(a[2]), "+r"(a[3]), "+r"(b[0]), "+r"(b[1]), "+r"(b[2]), "+r"(b[3])
                 :
                 : "memory");
}
__device__ __forceinline__ void wait_ld_three(uint32_t (&r)[8], uint32_t (&a)[4], uint32_t (&b)[4]) {
    asm volatile("tcgen05.wait::ld.sync.aligned;"
                 : "+r"(r[0]), "+r"(r[1]), "+r"(r[2]), "+r"(r[3]), "+r"(r[4]), "+r"(r[5]), "+r"(r[6]), "+r"(r[7]),
                   "+r"(a[0]), "+r"(a[1]), "+r"(a[2]), "+r"(a[3]), "+r"(b[0]), "+r"(b[1]), "+r"(b[2]), "+r"(b[3])
                 :
                 : "memory");
}
// Wait for every outstanding tcgen05.ld of the thread: two transposed blocks, or those and a block of 16 words.
__device__ __forceinline__ void wait_ld_pair(uint32_t (&a)[8], uint32_t (&b)[8]) {
    asm volatile("tcgen05.wait::ld.sync.aligned;"
                 : "+r"(a[0]), "+r"(a[1]), "+r"(a[2]), "+r"(a[3]), "+r"(a[4]), "+r"(a[5]), "+r"(a[6]), "+r"(a[7]),
                   "+r"(b[0]), "+r"(b[1]), "+r"(b[2]), "+r"(b[3]), "+r"(b[4]), "+r"(b[5]), "+r"(b[6]), "+r"(b[7])
                 :
                 : "memory");
}
__device__ __forceinline__ void wait_ld_three(uint32_t (&r)[16], uint32_t (&a)[8], uint32_t (&b)[8]) {
    asm volatile("tcgen05.wait::ld.sync.aligned;"
                 : "+r"(r[0]), "+r"(r[1]), "+r"(r[2]), "+r"(r[3]), "+r"(r[4]), "+r"(r[5]), "+r"(r[6]), "+r"(r[7]),
                   "+r"(r[8]), "+r"(r[9]), "+r"(r[10]), "+r"(r[11]), "+r"(r[12]), "+r"(r[13]), "+r"(r[14]), "+r"(r[15]),
                   "+r"(a[0]), "+r"(a[1]), "+r"(a[2]), "+r"(a[3]), "+r"(a[4]), "+r"(a[5]), "+r"(a[6]), "+r"(a[7]),
                   "+r"(b[0]), "+r"(b[1]), "+r"(b[2]), "+r"(b[3]), "+r"(b[4]), "+r"(b[5]), "+r"(b[6]), "+r"(b[7])
                 :
                 : "memory");
}

// kU reals <-> 32-bit words (register renames, no instructions).
template <int N>
__device__ __forceinline__ void pack(const double (&v)[N], uint32_t (&w)[2 * N]) {
#pragma unroll
    for (int u = 0; u < N; ++u) {
        w[2 * u] = (uint32_t)__double2loint(v[u]);
        w[2 * u + 1] = (uint32_t)__double2hiint(v[u]);
    }
}
template <int N>
__device__ __forceinline__ void unpack(const uint32_t (&w)[2 * N], double (&v)[N]) {
#pragma unroll
    for (int u = 0; u < N; ++u) v[u] = __hiloint2double((int)w[2 * u + 1], (int)w[2 * u]);
}
template <int N>
__device__ __forceinline__ void pack(const float (&v)[N], uint32_t (&w)[N]) {
#pragma unroll
    for (int u = 0; u < N; ++u) w[u] = __float_as_uint(v[u]);
}
template <int N>
__device__ __forceinline__ void unpack(const uint32_t (&w)[N], float (&v)[N]) {
#pragma unroll
    for (int u = 0; u < N; ++u) v[u] = __uint_as_float(w[u]);
}

}  // namespace tmem
}  // namespace cmb
