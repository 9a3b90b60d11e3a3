// placeholder until the wavefront integrator lands
#include "ctx.h"
namespace bdpt {
int render_impl(bdpt_ctx* ctx, const bdpt_camera*, const bdpt_render_opts*) { ctx->err = "render not built"; return BDPT_ERR_STATE; }
void render_free(bdpt_ctx*) {}
}
