// gemv_sym_int4.cu -- instantiates the dequant-GEMV kernel (gemv_kernel.cuh) for sym_int4.
#include "gemv_kernel.cuh"
namespace lowbit {
LOWBIT_DEFINE_GEMV_QT(sym_int4, QT_SYM_INT4)
}
