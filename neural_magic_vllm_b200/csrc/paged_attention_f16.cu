// fp16 instantiations of the tensor-core paged attention kernel.
#include "paged_attention_mma.cuh"
namespace b200 {
template int launch_paged_attn_mma<__half>(const AttnArgs&, int, dim3, size_t, cudaStream_t);
}
