// gemv_fp8_e5m2.cu -- instantiates the dequant-GEMV kernel (gemv_kernel.cuh) for fp8_e5m2.
#include "gemv_kernel.cuh"
namespace lowbit {
LOWBIT_DEFINE_GEMV_QT(fp8_e5m2, QT_FP8_E5M2)
}
