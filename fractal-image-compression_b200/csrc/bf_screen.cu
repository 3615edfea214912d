// placeholder until the tcgen05 screen lands
#include "fic_common.cuh"
bool bf_screen_applicable(int, int) { return false; }
int bf_screen_encode8(fic_handle*, const uint8_t*, int, int, int, int, int32_t*, double*, double*, unsigned, cudaStream_t)
{
    fic_set_error("screen path not built");
    return FIC_EINVAL;
}
