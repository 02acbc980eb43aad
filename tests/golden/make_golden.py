"""Generates tests/golden/*.npz by running the REAL reference (TarasKutsyk/Count_PIPNet, imported from /root/reference)
on the seeded synthetic weights/inputs. Run in the dev container only:  python tests/golden/make_golden.py

/root/reference is read-only and does not exist on the GPU box, so only the outputs travel (small .npz files). The
weights are not stored: count_pipnet_b200/synth.py regenerates them bit-identically from (key, shape, seed).
`matplotlib` is stubbed because pipnet/count_pipnet.py:10 imports it for a plot helper only.
"""
import os
import sys
import types

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import cases  # noqa: E402


def import_reference():
    for name in ("matplotlib", "matplotlib.pyplot"):
        if name not in sys.modules:
            sys.modules[name] = types.ModuleType(name)
    sys.modules["matplotlib"].pyplot = sys.modules["matplotlib.pyplot"]
    sys.path.insert(0, "/root/reference")
    import pipnet.pipnet as ref_pipnet
    import pipnet.count_pipnet as ref_count
    assert ref_pipnet.__file__.startswith("/root/reference"), ref_pipnet.__file__
    return ref_pipnet, ref_count


def summarise(proto: torch.Tensor) -> dict:
    """Small, checkable views of the [B,P,H,W] map: per-(b,p) max + flat argmax, per-(b,p) sum, a strided sample."""
    b, p, h, w = proto.shape
    flat = proto.reshape(b, p, h * w)
    mx, am = flat.max(dim=2)
    ps, hs = max(1, p // 16), max(1, h // 6)
    return dict(map_max=mx.numpy(), map_argmax=am.numpy().astype(np.int32), map_sum=flat.sum(2).numpy(),
                map_sample=proto[:, ::ps, ::hs, ::hs].contiguous().numpy(), map_shape=np.array(proto.shape))


def main():
    ref_pipnet, ref_count = import_reference()
    import oracle
    torch.set_num_threads(os.cpu_count() or 1)
    for name, case in cases.CASES.items():
        model = cases.build(case, ref_pipnet, ref_count)
        x = cases.inputs(case)
        torch.manual_seed(cases.GUMBEL_SEED)  # the only RNG use in an eval forward is F.gumbel_softmax's noise
        with torch.no_grad():
            proto, pooled, out = model(x, inference=case["inference"])
            feats = model._net(x)
        rec = dict(pooled=pooled.numpy(), out=out.numpy(), pred=out.argmax(1).numpy().astype(np.int32),
                   top_proto=pooled.argmax(1).numpy().astype(np.int32),
                   feat_absmax=np.float32(feats.abs().max().item()), feat_shape=np.array(feats.shape))
        rec.update(summarise(proto))
        # pin the oracle against the reference right here as well
        noise = cases.gumbel_noise(proto.shape) if case["args"].get("activation") == "gumbel_softmax" else None
        sd = {k: v.detach() for k, v in model.state_dict().items()}
        o_proto, o_pooled, o_out = oracle.forward(sd, cases.oracle_cfg(case), x, inference=case["inference"],
                                                  gumbel_noise=noise)
        d_pool = (o_pooled - pooled).abs().max().item()
        d_out = (o_out - out).abs().max().item() / max(out.abs().max().item(), 1e-6)
        d_map = (o_proto - proto).abs().max().item()
        print(f"{name:45s} feat|max|={rec['feat_absmax']:.2f} grid={tuple(proto.shape[2:])} "
              f"oracle-vs-reference: pooled {d_pool:.1e} out(rel) {d_out:.1e} map {d_map:.1e}")
        assert d_pool <= 2e-6 * max(1.0, pooled.abs().max().item()) and d_out <= 2e-6 and d_map <= 2e-6, name
        np.savez_compressed(os.path.join(HERE, name + ".npz"), **rec)


if __name__ == "__main__":
    main()
