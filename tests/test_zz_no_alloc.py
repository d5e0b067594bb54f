"""pg_spawn_epilogue_forward_static (include/pings_b200.h): the allocation-free variant of the spawn epilogue, whose
look-back scratch is a file-scope `static __device__` array inside the library.  No counterpart in the reference;
the oracle here is the existing entry point pg_spawn_epilogue_forward
on the same inputs: integer outputs bit-exact, float outputs bit-exact too (same kernel, same arithmetic).

Collected last on purpose (file name).
"""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu


def _inputs(n, K, seed, dev):
    g = torch.Generator().manual_seed(seed)
    r = lambda *s: torch.randn(*s, generator=g)
    N = n + 7
    c = dict(position=r(N, 3) * 5, orientation=torch.nn.functional.normalize(r(N, 4), dim=1), color=torch.rand(N, 3, generator=g),
             cam_origin=r(3), sidx=torch.randperm(N, generator=g)[:n].to(torch.int64),
             free_mask=(torch.rand(N, generator=g) < 0.3).to(torch.uint8))
    raws = dict(xyz_raw=r(n, 3 * K), rot_raw=r(n, 4 * K), scale_raw=r(n, 3 * K) * 0.5,
                alpha_raw=r(n, K), color_raw=r(n, 3 * K))
    return {k: v.to(dev).contiguous() for k, v in c.items()}, {k: v.to(dev).contiguous() for k, v in raws.items()}


def _run(static, n, K, c, raws, dev):
    from pings_b200 import _lib as L
    lib = L.lib()
    NK = n * K
    a = L.PgSpawnArgs()
    a.n, a.K = n, K
    a.sidx, a.position, a.orientation = c["sidx"].data_ptr(), c["position"].data_ptr(), c["orientation"].data_ptr()
    a.color, a.cam_origin = c["color"].data_ptr(), c["cam_origin"].data_ptr()
    a.xyz_raw, a.rot_raw, a.scale_raw = raws["xyz_raw"].data_ptr(), raws["rot_raw"].data_ptr(), raws["scale_raw"].data_ptr()
    a.alpha_raw, a.color_raw = raws["alpha_raw"].data_ptr(), raws["color_raw"].data_ptr()
    a.disp_range, a.unit_scale, a.max_scale, a.z_far, a.scale_filter = 0.5, 0.05, 0.4, 60.0, 0.3
    a.surfel, a.residual, a.alpha_filter, a.xy_only = 1, 1, 1, 0
    f32 = dict(dtype=torch.float32, device=dev)
    out = dict(xyz=torch.zeros(NK, 3, **f32), scale=torch.zeros(NK, 3, **f32), rot=torch.zeros(NK, 4, **f32),
               alpha=torch.zeros(NK, 1, **f32), color=torch.zeros(NK, 3, **f32), alpha_all=torch.zeros(NK, 1, **f32),
               cpos=torch.zeros(NK, dtype=torch.int32, device=dev), free_out=torch.zeros(NK, dtype=torch.uint8, device=dev))
    o = L.PgSpawnOut()
    o.xyz, o.scale, o.rot, o.alpha, o.color = (out[k].data_ptr() for k in ("xyz", "scale", "rot", "alpha", "color"))
    o.alpha_all, o.cpos, o.free_out, o.free_mask = (out["alpha_all"].data_ptr(), out["cpos"].data_ptr(),
                                                    out["free_out"].data_ptr(), c["free_mask"].data_ptr())
    st = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
    if static:
        cn = torch.full((2,), -1, dtype=torch.int32, device=dev)
        o.state = None
        L.check(lib.pg_spawn_epilogue_forward_static(C.byref(a), C.byref(o), cn.data_ptr(), st), "static spawn epilogue")
    else:
        state = torch.empty(lib.pg_spawn_state_words(NK), dtype=torch.int32, device=dev)
        o.state = state.data_ptr()
        L.check(lib.pg_spawn_epilogue_forward(C.byref(a), C.byref(o), st), "spawn epilogue")
        ci = lib.pg_spawn_count_index(NK)
        cn = state[ci:ci + 2].clone()
    torch.cuda.synchronize(dev)
    out["count_nan"] = cn
    return out


def test_static_scratch_entry_point_matches_allocating_one_twice_in_a_row():
    dev = torch.device("cuda:0")
    # two calls in a row with DIFFERENT sizes, the larger first: the second call must not see the first one's
    # look-back words, ticket or count in the shared array
    for n, K, seed in ((40000, 8, 1), (1237, 3, 2), (40000, 8, 3)):
        c, raws = _inputs(n, K, seed, dev)
        ref = _run(False, n, K, c, raws, dev)
        got = _run(True, n, K, c, raws, dev)
        M = int(ref["count_nan"][0])
        assert 0 < M < n * K and int(ref["count_nan"][1]) == 0
        assert torch.equal(got["count_nan"], ref["count_nan"])
        assert torch.equal(got["cpos"], ref["cpos"])
        assert torch.equal(got["free_out"][:M], ref["free_out"][:M])
        torch.testing.assert_close(got["alpha_all"], ref["alpha_all"], rtol=0, atol=0)
        for k in ("xyz", "scale", "rot", "alpha", "color"):
            torch.testing.assert_close(got[k][:M], ref[k][:M], rtol=0, atol=0)


def test_static_scratch_entry_point_rejects_oversize_and_allocates_nothing():
    from pings_b200 import _lib as L
    lib = L.lib()
    dev = torch.device("cuda:0")
    n, K = 3000, 8
    c, raws = _inputs(n, K, 5, dev)
    _run(True, n, K, c, raws, dev)                       # set-up (symbol address looked up) outside the measurement
    torch.cuda.synchronize(dev)
    free0, _ = torch.cuda.mem_get_info(dev)
    a = L.PgSpawnArgs()
    a.n, a.K = (1 << 24) // 8 + 1, 8                      # n*K just above PG_SPAWN_STATIC_MAX_NK: rejected before any pointer is used
    for f in ("position", "orientation", "xyz_raw", "rot_raw", "scale_raw", "alpha_raw", "color_raw"):
        setattr(a, f, raws["xyz_raw"].data_ptr())
    o = L.PgSpawnOut()
    for f in ("xyz", "scale", "rot", "alpha", "color", "alpha_all", "cpos"):
        setattr(o, f, raws["xyz_raw"].data_ptr())
    cn = torch.zeros(2, dtype=torch.int32, device=dev)
    rc = lib.pg_spawn_epilogue_forward_static(C.byref(a), C.byref(o), cn.data_ptr(), None)
    assert rc == 3      # PG_ERR_UNSUPPORTED
    assert b"PG_SPAWN_STATIC_MAX_NK" in lib.pg_last_error()
    free1, _ = torch.cuda.mem_get_info(dev)
    assert free1 == free0                                 # no device memory taken by a call
