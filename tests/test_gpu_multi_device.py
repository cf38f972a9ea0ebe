"""One process, several GPUs: INTEGRATION.md promises that a caller may drive each device from its own host
thread (cudaSetDevice before the call).  The launch state the library caches (shared-memory opt-in of each
kernel instantiation, SM count, the persistent kernels' item-counter pools) is per DEVICE; round 1 kept it in
process-wide statics, so the first launch on a second device failed.  Skipped when fewer than two devices
are visible (the round-end GPU tests run on one); run with `gpurun --gpus 2`."""
import threading

import pytest
import torch

pytestmark = pytest.mark.gpu


def _work(dev, results, barrier):
    try:
        from flash_attention_hip_b200 import api
        torch.cuda.set_device(dev)
        g = torch.Generator(device=f"cuda:{dev}").manual_seed(7)       # the same inputs on every device
        mk = lambda *s: torch.randn(s, generator=g, device=f"cuda:{dev}", dtype=torch.float32).bfloat16()
        got = {}
        # persistent launches (items > SMs) and one-item-per-CTA launches, dense and packed, several instantiations
        for name, (b, s, h, hk, d, p) in {"persistent_d128_drop": (8, 1024, 16, 4, 128, 0.1),
                                          "small_d64": (1, 256, 4, 4, 64, 0.0),
                                          "persistent_d64": (16, 1024, 16, 16, 64, 0.0)}.items():
            q, k, v, do = mk(b, s, h, d), mk(b, s, hk, d), mk(b, s, hk, d), mk(b, s, h, d)
            barrier.wait()                                           # launch on all devices at the same time
            for _ in range(3):
                out, lse, _, _ = api.flash_attn_fwd(q, k, v, p_dropout=p, is_causal=True, seed=5, offset=1)
                assert api.last_path() == api.PATH_TCGEN05
                dq, dk, dv = api.flash_attn_bwd(do, q, k, v, out, lse, p_dropout=p, is_causal=True, seed=5, offset=1,
                                                deterministic=True)
                assert api.last_path() == api.PATH_TCGEN05
            torch.cuda.synchronize(dev)
            got[name] = [t.cpu() for t in (out, lse, dq, dk, dv)]
        results[dev] = got
    except Exception as exc:  # noqa: BLE001 -- reported by the main thread
        results[dev] = exc
        try:
            barrier.abort()
        except Exception:
            pass


def test_two_devices_from_two_host_threads_of_one_process():
    n = torch.cuda.device_count()
    if n < 2:
        pytest.skip("needs two visible devices")
    n = min(n, 4)
    results, barrier = {}, threading.Barrier(n)
    threads = [threading.Thread(target=_work, args=(dev, results, barrier)) for dev in range(n)]
    for t in threads:
        t.start()
    for t in threads:
        t.join()
    for dev in range(n):
        assert not isinstance(results.get(dev), Exception), f"device {dev}: {results.get(dev)!r}"
    # same inputs, deterministic kernels: every device returns the same bits
    for name in results[0]:
        for a, b in zip(results[0][name], results[1][name]):
            assert torch.equal(a, b), name


def test_second_device_after_first_in_one_thread():
    """The simplest form of the round-1 defect: device 0 first, then device 1, same thread."""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two visible devices")
    from flash_attention_hip_b200 import api
    outs = []
    for dev in (0, 1, 0):
        with torch.cuda.device(dev):
            g = torch.Generator(device=f"cuda:{dev}").manual_seed(3)
            q, k, v = (torch.randn((4, 2048, 16, 128), generator=g, device=f"cuda:{dev}").bfloat16() for _ in range(3))
            out, lse, _, _ = api.flash_attn_fwd(q, k, v, is_causal=True)
            torch.cuda.synchronize(dev)
            outs.append(out.cpu())
    assert torch.equal(outs[0], outs[1]) and torch.equal(outs[0], outs[2])
