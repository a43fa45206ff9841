// gpb_mh.cu -- placeholder translation unit; filled in by the MH / matrix-solve milestone.
#include "gpb_internal.h"
extern "C" {
int gpb_trsm_lower(gpb_ctx*, int, int, const double*, double*) { return GPB_E_UNSUPPORTED; }
int gpb_trsm_upper(gpb_ctx*, int, int, const double*, double*) { return GPB_E_UNSUPPORTED; }
int gpb_mh_run(gpb_ctx*, const gpb_dataset*, int, int, const gpb_term*, int, const double*, int,
               const double*, const gpb_mh_opts*, const double*, const double*, double*, double*,
               int*, int*) { return GPB_E_UNSUPPORTED; }
int gpb_philox_streams(unsigned long long, int, int, int, int, double*, double*) { return GPB_E_UNSUPPORTED; }
}
