"""Threading contract of the C ABI (include/rt_abi.h "Threading"): calls on one scene handle from several OS threads are
serialised inside the library and give the single-threaded results; different handles run concurrently; a handle destroyed
while a call is in flight stays valid for that call; rt_last_error is thread-local.  (ctypes releases the GIL around foreign
calls, as a Haskell `safe` import releases the capability.)"""
import ctypes as C
import threading

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

W, H = 60, 40


def test_concurrent_calls_on_one_handle(gpu):
    ds = gpu.HostScene("cornell").upload(W, H)
    serial = [ds.render(W, H, 6, seed=100 + i)[0] for i in range(6)]
    out = [None] * 6
    errs = []

    def work(i):
        try:
            out[i] = ds.render(W, H, 6, seed=100 + i)[0]
        except Exception as e:  # noqa: BLE001
            errs.append(e)
    th = [threading.Thread(target=work, args=(i,)) for i in range(6)]
    [t.start() for t in th]; [t.join() for t in th]
    assert not errs, errs
    for a, b in zip(out, serial):
        assert np.allclose(a, b, rtol=1e-12, atol=1e-12)
    ds.destroy()


def test_two_handles_and_thread_local_errors(gpu):
    a = gpu.HostScene("cornell").upload(W, H)
    b = gpu.HostScene("two_spheres").upload(W, H)
    ref_a, ref_b = a.render(W, H, 8, seed=3)[0], b.render(W, H, 8, seed=3)[0]
    res = {}

    def good(name, ds):
        res[name] = ds.render(W, H, 8, seed=3)[0]

    def bad():
        try:
            a.render(W, H, 1, region=(10, 5, 0, W))
        except RuntimeError as e:
            res["bad"] = str(e)
        res["bad_last"] = gpu.core().rt_last_error().decode()
    th = [threading.Thread(target=good, args=("a", a)), threading.Thread(target=good, args=("b", b)), threading.Thread(target=bad)]
    [t.start() for t in th]; [t.join() for t in th]
    assert np.allclose(res["a"], ref_a, rtol=1e-12, atol=1e-12) and np.allclose(res["b"], ref_b, rtol=1e-12, atol=1e-12, equal_nan=True)
    assert "bad region" in res["bad"] and "bad region" in res["bad_last"]
    a.destroy(); b.destroy()


def test_destroy_while_rendering(gpu):
    hs = gpu.HostScene("cornell")
    ds = hs.upload(200, 200)
    ref = ds.render(200, 200, 64, seed=9)[0]
    got = {}
    started = threading.Event()

    def work():
        started.set()
        try:
            got["img"] = ds.render(200, 200, 64, seed=9)[0]
        except RuntimeError as e:  # destroy won the race before the call looked the handle up
            got["err"] = str(e)
    t = threading.Thread(target=work)
    t.start(); started.wait()
    sid = C.c_uint64(ds.id.value)
    assert gpu.core().rt_scene_destroy(sid) == 0  # races with the render: before it, or during it (then it waits for the call)
    t.join()
    ds.id = None
    if "img" in got:
        assert np.allclose(got["img"], ref, rtol=1e-12, atol=1e-12)
    else:
        assert "unknown scene handle" in got["err"]
    assert gpu.core().rt_scene_destroy(sid) == -4  # RT_ERR_HANDLE: the handle is gone either way
    p = gpu.DeviceScene.params(8, 8, 1)
    out = np.zeros((8, 8, 3))
    assert gpu.core().rt_render(sid, C.byref(p), out.ctypes.data_as(C.c_void_p), None) == -4
    assert b"unknown scene handle" in gpu.core().rt_last_error()
