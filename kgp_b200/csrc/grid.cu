// KISS-GP / MSGP (infGrid) path -- placeholder until the grid kernels land.
#include "ctx.h"
namespace kgp {
int grid_eval(kgp_ctx* c, int, int, double*, double*) { return c->fail(KGP_BAD_ARG, "infGrid path not built yet"); }
int grid_dlik(kgp_ctx* c, int, double*, bool) { return c->fail(KGP_BAD_ARG, "infGrid path not built yet"); }
int grid_alpha(kgp_ctx* c, double*) { return c->fail(KGP_BAD_ARG, "infGrid path not built yet"); }
int grid_predict(kgp_ctx* c, const double*, long, double*, double*, double*, double*) { return c->fail(KGP_BAD_ARG, "infGrid path not built yet"); }
int grid_create(kgp_ctx* c, int, int, double) { return c->fail(KGP_BAD_ARG, "infGrid path not built yet"); }
void grid_free(kgp_ctx*) {}
}
