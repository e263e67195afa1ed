// prefill_tcgen05.cu — placeholder until the tcgen05/TMEM kernel lands.
#include "common.cuh"
namespace mvb200 {
struct PrefillArgs;
int prefill_tcgen05_dispatch(const PrefillArgs&, cudaStream_t) { return -1; }
}  // namespace mvb200
