"""CPU suite: the ticket schedule of the pipelined single-launch kernels (gpu-fhe_b200/csrc/pipe_sched.h), decoded by the
very function the kernels run.  Deadlock freedom of the in-kernel waits rests on two properties checked here for many
shapes: every (pass, polynomial, tile) is drawn exactly once, and every tile a ticket waits for has a smaller ticket."""
import ctypes

import pytest


def schedule(passes, tiles, npoly, resident):
    import arithmetic  # noqa: F401
    from gpu_fhe_b200 import _capi
    n = (ctypes.c_uint32 * 3)(*(list(tiles) + [0] * (3 - len(tiles))))
    lag = ctypes.c_uint32()
    total = _capi.lib.gfhe_debug_pipe_schedule(passes, n, npoly, resident, None, 0, ctypes.byref(lag))
    assert total == npoly * sum(tiles)
    out = (ctypes.c_uint32 * (3 * total))()
    assert _capi.lib.gfhe_debug_pipe_schedule(passes, n, npoly, resident, out, total, None) == total
    return [(out[3 * t], out[3 * t + 1], out[3 * t + 2]) for t in range(total)], lag.value


@pytest.mark.parametrize("passes,tiles", [(2, (16, 16)), (3, (32, 16, 16)), (3, (16, 16, 16)), (2, (2, 2)), (3, (4, 2, 2)), (2, (256, 256)), (2, (1, 1))])
@pytest.mark.parametrize("npoly", [1, 2, 3, 7, 64, 513])
@pytest.mark.parametrize("resident", [1, 37, 296, 592])
def test_ticket_order(built_lib, passes, tiles, npoly, resident):
    if npoly * sum(tiles) > 150000:
        pytest.skip("large")
    tickets, lag = schedule(passes, tiles, npoly, resident)
    assert 1 <= lag <= npoly
    seen = {}
    last_of = {}                      # (pass, poly) -> largest ticket among its tiles
    for t, (k, p, w) in enumerate(tickets):
        assert k < passes and p < npoly and w < tiles[k], (t, k, p, w)
        assert (k, p, w) not in seen
        seen[(k, p, w)] = t
        last_of[(k, p)] = t
    assert len(seen) == npoly * sum(tiles)            # every tile of every pass exactly once
    for (k, p, w), t in seen.items():
        if k > 0:                                     # all producers (pass k-1 of the same polynomial) were drawn earlier
            assert last_of[(k - 1, p)] < t
    # the lag keeps a polynomial's consumers `lag` steps behind its producers: at least (lag - 1) steps' worth of tickets
    # (a step draws at least the first pass's tiles) lie between the last producer and the first consumer
    for p in range(npoly):
        first_consumer = min(seen[(1, p, w)] for w in range(tiles[1]))
        assert first_consumer - last_of[(0, p)] >= (lag - 1) * min(tiles[0], sum(tiles)) * (1 if p + lag <= npoly else 0)
    z = sum(tiles)
    assert lag == min(npoly, max(1, -(-(resident + resident // 4) // z)))
