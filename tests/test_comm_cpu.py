"""Host-side behaviour of the communicator module that needs no GPU."""
import pytest
import torch


def test_unsupported_world_size_is_a_clear_error():
    from ddp_b200.parallel.comm import SUPPORTED_WORLD_SIZES, SymmetricComm
    assert SUPPORTED_WORLD_SIZES == tuple(range(1, 9))  # any size up to one NVSwitch node of 8
    for rank, world in ((0, 9), (0, 16), (2, 2), (-1, 2), (0, 0)):
        with pytest.raises(ValueError, match="world sizes"):
            SymmetricComm(rank, world, 1024)


def test_conv3x3_wgrad_image_dest_is_a_bijection_onto_the_bucket():
    """The index map of the fused conv2 update: every weight and bias element of net.2 is hit exactly
    once, the 7 spare bias columns per output channel are dropped."""
    from ddp_b200.parallel.comm import CONV3X3_WGRAD_IMAGE, conv3x3_wgrad_image_dest
    dest = conv3x3_wgrad_image_dest()
    assert dest.numel() == CONV3X3_WGRAD_IMAGE == 64 * (9 * 32 + 8)
    kept = dest[dest >= 0]
    assert kept.numel() == 64 * 32 * 9 + 64
    assert torch.equal(torch.sort(kept).values, torch.arange(64 * 32 * 9 + 64))
    assert int((dest < 0).sum()) == 64 * 7
    # spot checks against the accumulator layout: channel co, row r, tap sx, input channel ci
    co, r, sx, ci = 5, 2, 1, 17
    col = (3 * 32 + 8) + 3 * 32 + sx * 32 + ci          # row 2 block starts after row 0 (104) and row 1 (96)
    assert int(dest[co * 296 + col]) == (co * 32 + ci) * 9 + r * 3 + sx
    assert int(dest[co * 296 + 96]) == 64 * 32 * 9 + co  # first bias column


def test_bucket_plan_cuts_inside_large_parameters():
    """DDP caps buckets at 25 MiB; a single 94 MB gradient (the widened CNN's fc1) must not become one
    94 MB exchange: the plan cuts through it, covers the arena exactly once and keeps 16-byte alignment."""
    from ddp_b200.parallel.bucketed import plan_buckets
    # arena order of the widened CNN (reverse registration): fc2.b, fc2.w, fc1.b, fc1.w, conv2.b, conv2.w, ...
    sizes = [10, 7500, 750, 32768 * 750, 128, 128 * 64 * 9, 64, 64 * 3 * 9]
    spans, off = [], 0
    for n in sizes:
        off = (off + 3) // 4 * 4
        spans.append((off, n))
        off += n
    cuts = plan_buckets(spans, 25.0)
    cap = 25 * (1 << 20) // 4
    assert cuts[0][0] == 0 and cuts[-1][1] == (off + 3) // 4 * 4
    assert all(a[1] == b[0] for a, b in zip(cuts, cuts[1:]))
    assert all((e - s) % 4 == 0 and 0 < e - s <= cap for s, e in cuts)
    assert len(cuts) >= 4                      # 94 MB / 25 MiB
    assert plan_buckets(spans, 1000.0) == [(0, cuts[-1][1])]
