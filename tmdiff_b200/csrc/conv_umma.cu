// placeholder until the tcgen05 kernel lands
#include "kernels.h"
namespace tmd {
struct UmmaPlan { int dummy; };
bool conv_umma_supported(const ConvOp&) { return false; }
UmmaPlan* conv_umma_plan_create(const ConvOp&) { return nullptr; }
void conv_umma_plan_destroy(UmmaPlan* p) { delete p; }
cudaError_t conv_umma_bf16(const ConvOp&, const UmmaPlan*, cudaStream_t) { return cudaErrorNotSupported; }
}
