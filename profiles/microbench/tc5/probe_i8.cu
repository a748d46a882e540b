// Does ptxas accept tcgen05.mma.kind::i8 for sm_100a?    Compile only:
//   nvcc -gencode arch=compute_100a,code=sm_100a -c probe_i8.cu -o /dev/null
#include <cstdint>
__global__ void probe(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc) {
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(1));
}
