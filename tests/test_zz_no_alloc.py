"""bdpt_film_to_rgb8_noalloc (include/bdpt.h): the tone-map through the library's static device scratch, for callers
that may not call cudaMalloc.  Checked through the C-ABI against the existing bdpt_film_to_rgb8 (byte-exact: it is
integer output of the same kernel), on calls in a row that share the scratch.  Collected last on purpose;
it stands alone: its own context, no state from the other tests."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def test_noalloc_tonemap_equals_the_allocating_one_on_calls_in_a_row(bdpt, scenes):
    v, m, mats, cam = scenes["cornell_spheres"]()
    with bdpt.Context(0) as ctx:
        ctx.upload(v, m, mats)
        # two frames of different size, largest first: the second call finds the first frame's bytes in the scratch
        for k, (W, H, spp) in enumerate([(40, 28, 4), (24, 24, 9), (40, 28, 4)]):
            ctx.render(bdpt.make_camera(W, H, spp, 6, **cam), seed=5 + k, bvh=bdpt.BVH_FAST, height=H)
            want = ctx.film_rgb8(spp)                      # existing entry point (cudaMalloc'ed staging)
            assert want.any() and want.shape == (H, W, 3)
            got = np.full((H, W, 3), 0xA5, np.uint8)
            ctx.film_rgb8_noalloc(spp, out=got)            # new entry point (static scratch)
            assert np.array_equal(got, want), "call %d differs" % k
            again = ctx.film_rgb8_noalloc(spp)             # immediately again: same answer from the shared scratch
            assert np.array_equal(again, want)
        # a second context on the same device shares the array (sequential calls: no race)
        with bdpt.Context(0) as other:
            other.film_clear(8, 8)
            assert not other.film_rgb8_noalloc(1).any()
        assert np.array_equal(ctx.film_rgb8_noalloc(4), want)


def test_noalloc_tonemap_refuses_films_beyond_the_static_scratch(bdpt):
    with bdpt.Context(0) as ctx:
        W, H = 4096, 2731                                   # 33,558,528 bytes > 32 MiB
        assert W * H * 3 > bdpt.NOALLOC_RGB8_MAX_BYTES
        ctx.film_clear(W, H)
        out = np.zeros((H, W, 3), np.uint8)
        rc = bdpt.lib().bdpt_film_to_rgb8_noalloc(ctx._h, 1, out.ctypes.data)
        assert rc == -1                                     # BDPT_ERR_ARG
        msg = bdpt.lib().bdpt_last_error(ctx._h).decode()
        assert "33558528" in msg and str(bdpt.NOALLOC_RGB8_MAX_BYTES) in msg
        assert not out.any()
        assert ctx.film_rgb8(1).shape == (H, W, 3)          # the allocating entry point still takes it
        # exactly at the limit is accepted: 4096 x 2730 x 3 = 33,546,240 <= 33,554,432
        ctx.film_clear(4096, 2730)
        assert not ctx.film_rgb8_noalloc(1).any()
    with bdpt.Context(0) as fresh:                          # no film yet: BDPT_ERR_STATE, not a fault
        byte = np.zeros(3, np.uint8)
        assert bdpt.lib().bdpt_film_to_rgb8_noalloc(fresh._h, 1, byte.ctypes.data) == -3
