// gemv_asym_int4.cu -- instantiates the dequant-GEMV kernel (gemv_kernel.cuh) for asym_int4.
#include "gemv_kernel.cuh"
namespace lowbit {
LOWBIT_DEFINE_GEMV_QT(asym_int4, QT_ASYM_INT4)
}
