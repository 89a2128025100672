// placeholder until the tcgen05 kernel lands (next commit)
#include "model.h"
namespace nb {
struct PackedBf16 { int dummy; };
void packed_free(PackedBf16* p) { delete p; }
int packed_build(nb_model* m, cudaStream_t) { set_error("bf16 path not built"); return NB_ERR_UNSUPPORTED; }
int tc_forward(const nb_model*, const float*, int, const float*, const float*, const float*, int, int64_t, int, int64_t, float*, void*, size_t, int, cudaStream_t) { set_error("bf16 path not built"); return NB_ERR_UNSUPPORTED; }
int tc_backward(const nb_model*, int64_t, float*, void*, size_t, float*, cudaStream_t) { set_error("bf16 path not built"); return NB_ERR_UNSUPPORTED; }
size_t tc_workspace_bytes(const nb_model*, int64_t, int) { return 0; }
}
