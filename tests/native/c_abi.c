#include "aqgpu.h"
int main(void) {
    aqgpu_agg_t a = {0};
    aqgpu_window_t w = {0};
    (void)a; (void)w;
    return aqgpu_abi_version() == AQGPU_ABI_VERSION ? 0 : 1;
}
