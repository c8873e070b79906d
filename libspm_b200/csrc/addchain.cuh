// addchain.cuh -- generated: s = a + b over NW 32-bit words with hardware carry (IADD3 / IADD3.X)
#pragma once
#include <stdint.h>

template <int NW> __device__ __forceinline__ void spmb_add_chain(uint32_t (&s)[NW], const uint32_t (&a)[NW], const uint32_t (&b)[NW]);

template <> __device__ __forceinline__ void spmb_add_chain<1>(uint32_t (&s)[1], const uint32_t (&a)[1], const uint32_t (&b)[1])
{
    asm("add.u32 %0, %1, %2;"
        : "=r"(s[0])
        : "r"(a[0]), "r"(b[0]));
}

template <> __device__ __forceinline__ void spmb_add_chain<2>(uint32_t (&s)[2], const uint32_t (&a)[2], const uint32_t (&b)[2])
{
    asm("add.cc.u32 %0, %2, %4;\n\taddc.u32 %1, %3, %5;"
        : "=&r"(s[0]), "=r"(s[1])
        : "r"(a[0]), "r"(a[1]), "r"(b[0]), "r"(b[1]));
}

template <> __device__ __forceinline__ void spmb_add_chain<3>(uint32_t (&s)[3], const uint32_t (&a)[3], const uint32_t (&b)[3])
{
    asm("add.cc.u32 %0, %3, %6;\n\taddc.cc.u32 %1, %4, %7;\n\taddc.u32 %2, %5, %8;"
        : "=&r"(s[0]), "=&r"(s[1]), "=r"(s[2])
        : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(b[0]), "r"(b[1]), "r"(b[2]));
}

template <> __device__ __forceinline__ void spmb_add_chain<4>(uint32_t (&s)[4], const uint32_t (&a)[4], const uint32_t (&b)[4])
{
    asm("add.cc.u32 %0, %4, %8;\n\taddc.cc.u32 %1, %5, %9;\n\taddc.cc.u32 %2, %6, %10;\n\taddc.u32 %3, %7, %11;"
        : "=&r"(s[0]), "=&r"(s[1]), "=&r"(s[2]), "=r"(s[3])
        : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b[0]), "r"(b[1]), "r"(b[2]), "r"(b[3]));
}

template <> __device__ __forceinline__ void spmb_add_chain<5>(uint32_t (&s)[5], const uint32_t (&a)[5], const uint32_t (&b)[5])
{
    asm("add.cc.u32 %0, %5, %10;\n\taddc.cc.u32 %1, %6, %11;\n\taddc.cc.u32 %2, %7, %12;\n\taddc.cc.u32 %3, %8, %13;\n\taddc.u32 %4, %9, %14;"
        : "=&r"(s[0]), "=&r"(s[1]), "=&r"(s[2]), "=&r"(s[3]), "=r"(s[4])
        : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(a[4]), "r"(b[0]), "r"(b[1]), "r"(b[2]), "r"(b[3]), "r"(b[4]));
}

template <> __device__ __forceinline__ void spmb_add_chain<6>(uint32_t (&s)[6], const uint32_t (&a)[6], const uint32_t (&b)[6])
{
    asm("add.cc.u32 %0, %6, %12;\n\taddc.cc.u32 %1, %7, %13;\n\taddc.cc.u32 %2, %8, %14;\n\taddc.cc.u32 %3, %9, %15;\n\taddc.cc.u32 %4, %10, %16;\n\taddc.u32 %5, %11, %17;"
        : "=&r"(s[0]), "=&r"(s[1]), "=&r"(s[2]), "=&r"(s[3]), "=&r"(s[4]), "=r"(s[5])
        : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(a[4]), "r"(a[5]), "r"(b[0]), "r"(b[1]), "r"(b[2]), "r"(b[3]), "r"(b[4]), "r"(b[5]));
}

template <> __device__ __forceinline__ void spmb_add_chain<7>(uint32_t (&s)[7], const uint32_t (&a)[7], const uint32_t (&b)[7])
{
    asm("add.cc.u32 %0, %7, %14;\n\taddc.cc.u32 %1, %8, %15;\n\taddc.cc.u32 %2, %9, %16;\n\taddc.cc.u32 %3, %10, %17;\n\taddc.cc.u32 %4, %11, %18;\n\taddc.cc.u32 %5, %12, %19;\n\taddc.u32 %6, %13, %20;"
        : "=&r"(s[0]), "=&r"(s[1]), "=&r"(s[2]), "=&r"(s[3]), "=&r"(s[4]), "=&r"(s[5]), "=r"(s[6])
        : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(a[4]), "r"(a[5]), "r"(a[6]), "r"(b[0]), "r"(b[1]), "r"(b[2]), "r"(b[3]), "r"(b[4]), "r"(b[5]), "r"(b[6]));
}

template <> __device__ __forceinline__ void spmb_add_chain<8>(uint32_t (&s)[8], const uint32_t (&a)[8], const uint32_t (&b)[8])
{
    asm("add.cc.u32 %0, %8, %16;\n\taddc.cc.u32 %1, %9, %17;\n\taddc.cc.u32 %2, %10, %18;\n\taddc.cc.u32 %3, %11, %19;\n\taddc.cc.u32 %4, %12, %20;\n\taddc.cc.u32 %5, %13, %21;\n\taddc.cc.u32 %6, %14, %22;\n\taddc.u32 %7, %15, %23;"
        : "=&r"(s[0]), "=&r"(s[1]), "=&r"(s[2]), "=&r"(s[3]), "=&r"(s[4]), "=&r"(s[5]), "=&r"(s[6]), "=r"(s[7])
        : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(a[4]), "r"(a[5]), "r"(a[6]), "r"(a[7]), "r"(b[0]), "r"(b[1]), "r"(b[2]), "r"(b[3]), "r"(b[4]), "r"(b[5]), "r"(b[6]), "r"(b[7]));
}
