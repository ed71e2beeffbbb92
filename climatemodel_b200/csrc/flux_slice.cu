// The flux kernels of one (precision, molecule count) as a translation unit of their own: flux_kernel<Real, MM, V> for
// every variant V the library ships takes most of the library's compile time, and the ten slices are independent,
// so build.py compiles them in parallel (nvcc -c -DCMB_SLICE_REAL=double -DCMB_SLICE_M=2 -DCMB_SLICE_NAME=launch_flux_slice_d2)
// and flux.cu (built with -DCMB_FLUX_SLICES) calls them through these entry points.  No relocatable device code:
// every kernel is launched from the unit that instantiates it.
#define CMB_FLUX_SLICE_ONLY
#include "flux.cu"

namespace cmb {
int CMB_SLICE_NAME(int variant, cmb200_plan *plan, const cmb200_flux_io *io, const double *Tint, cudaStream_t s) {
    return launch_flux_variant<CMB_SLICE_REAL, CMB_SLICE_M>(variant, plan, io, Tint, s);
}
}  // namespace cmb
