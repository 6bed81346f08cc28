"""CPU tests of the weight preparation / layer-program builder (host logic)."""
import numpy as np
import torch

from oracle import resnet_oracle as ro
from tests import util


def test_program_shapes_and_flops():
    p50 = util.program_for("resnet50")
    assert sum(1 for l in p50.layers if l.op == 0) == 53 and p50.layers[-1].op == 2
    assert abs(p50.flops_per_image() / 1e9 - 8.178) < 1e-2
    p18 = util.program_for("resnet18")
    assert sum(1 for l in p18.layers if l.op == 0) == 20
    assert abs(p18.flops_per_image() / 1e9 - 3.628) < 1e-2
    p4 = util.program_for("resnet50", 4, 256)
    assert abs(p4.flops_per_image() / 1e9 - 10.78) < 1e-2 and p4.in_c == 4


def test_blob_alignment_and_bounds():
    p = util.program_for("resnet18")
    for l in p.layers:
        if l.op == 1:
            continue
        assert l.w_off % 256 == 0 and l.b_off % 256 == 0
        n = l.cout * l.kh * l.kw * l.cin * 2
        assert l.w_off + n <= len(p.blob) and l.b_off + l.cout * 4 <= len(p.blob)


def test_fold_matches_oracle_fold():
    """two independent BN folds (product: models.fold_conv_bn, oracle: resnet_oracle.fold_bn)."""
    from ai4e_b200.models import fold_conv_bn
    m = util.oracle_model("resnet18")
    sd = m.state_dict()
    for conv, bn, mod_c, mod_b in [("conv1", "bn1", m.conv1, m.bn1),
                                   ("layer2.0.downsample.0", "layer2.0.downsample.1", m.layer2[0].downsample[0],
                                    m.layer2[0].downsample[1]),
                                   ("layer4.1.conv2", "layer4.1.bn2", m.layer4[1].conv2, m.layer4[1].bn2)]:
        w, b = fold_conv_bn(sd, conv, bn)
        wo, bo = ro.fold_bn(mod_c, mod_b)
        assert torch.equal(w, wo.permute(0, 2, 3, 1).contiguous()) and torch.equal(b, bo)


def test_slot_liveness():
    """no layer overwrites a slot that a later layer still reads (simulated execution)."""
    for arch in ("resnet18", "resnet50"):
        for keep in (False, True):
            p = util.program_for(arch, keep_stages=keep)
            version = {0: 0}
            writes = 0
            producer_of = {}
            for i, l in enumerate(p.layers):
                for s in (l.in_buf, l.res_buf):
                    if s >= 0:
                        assert s in version, f"layer {i} reads unwritten slot {s}"
                if l.op != 2:
                    assert l.out_buf not in (l.in_buf, l.res_buf) and l.out_buf != 0
                    writes += 1
                    version[l.out_buf] = writes
                    producer_of[l.out_buf] = i
            # residual sources must be the block input or its downsample: produced before conv1 of the block
            if keep:
                slots = [v[0] for v in p.stage_bufs.values()]
                assert len(set(slots)) == len(slots)
                for name, (slot, *_rest) in p.stage_bufs.items():
                    last_writer = max(i for i, l in enumerate(p.layers) if l.op != 2 and l.out_buf == slot)
                    stage_end = {"stem": 0, "pool": 1}.get(name)
                    if stage_end is not None:
                        assert last_writer == stage_end


def test_stem_padding_channel_is_zero():
    p = util.program_for("resnet18")
    stem = p.layers[0]
    w = np.frombuffer(p.blob, np.uint16, stem.cout * 49 * 4, stem.w_off).reshape(stem.cout, 7, 7, 4)
    assert stem.cin == 4 and (w[..., 3] == 0).all() and (w[..., :3] != 0).any()
