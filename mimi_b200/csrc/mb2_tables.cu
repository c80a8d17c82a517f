// Device-side regeneration of the per-quadrature-point tables (placeholder until the MFEM-order
// NURBS evaluation lands; see mb2_generate_patch_tables in include/mimi_b200.h).
#include "mb2_ctx.h"
namespace mb2 {
int generate_patch_tables(mb2_ctx* c, int, const int32_t*, const double* const*, const int32_t*, const double*,
                          const double*, int) {
  c->err = "mb2_generate_patch_tables: not implemented yet";
  return MB2_ERR_UNSUPPORTED;
}
}  // namespace mb2
