// gemv_fp8_e4m3.cu -- instantiates the dequant-GEMV kernel (gemv_kernel.cuh) for fp8_e4m3.
#include "gemv_kernel.cuh"
namespace lowbit {
LOWBIT_DEFINE_GEMV_QT(fp8_e4m3, QT_FP8_E4M3)
}
