// gemv_nf4.cu -- instantiates the dequant-GEMV kernel (gemv_kernel.cuh) for nf4.
#include "gemv_kernel.cuh"
namespace lowbit {
LOWBIT_DEFINE_GEMV_QT(nf4, QT_NF4)
}
