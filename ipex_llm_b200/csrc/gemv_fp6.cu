// gemv_fp6.cu -- instantiates the dequant-GEMV kernel (gemv_kernel.cuh) for fp6.
#include "gemv_kernel.cuh"
namespace lowbit {
LOWBIT_DEFINE_GEMV_QT(fp6, QT_FP6)
}
