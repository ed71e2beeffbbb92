// Main translation unit of libclimatemodel_b200.so: the kernels are launched across what would otherwise be file
// boundaries, and one TU avoids relocatable device code.  The flux kernel instantiations alone live in translation
// units of their own (flux_slice.cu, one per precision and molecule count; build.py compiles them in parallel).
#include "plan.cu"
#include "tables.cu"
#include "flux.cu"
#include "noalloc.cu"
#include "transmission.cu"
#include "peaks.cu"
#include "lines.cu"
#include "convect.cu"
#include "stepper.cu"
#include "grey.cu"
