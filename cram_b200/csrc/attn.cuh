// Internal declarations shared by the attention translation units.
#pragma once
#include "common.cuh"

namespace cram {

int launch_attn_generic(const cram_attn_prefill_params& p, cudaStream_t stream);
// tcgen05 / TMEM / TMA kernel: bf16 or fp16, head_dim 64 or 128.
bool attn_sm100_supported(const cram_attn_prefill_params& p, const char** why);
int launch_attn_sm100(const cram_attn_prefill_params& p, cudaStream_t stream);

}  // namespace cram
