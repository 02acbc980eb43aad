"""Prints the parity table (every case x precision mode) of the fused CUDA forward against the reference golden
vectors. Run on the GPU box:  python tests/parity_report.py > gpurun_out/parity_report.txt"""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, HERE)

import cases  # noqa: E402
from count_pipnet_b200.pipnet import count_pipnet as my_count  # noqa: E402
from count_pipnet_b200.pipnet import pipnet as my_pipnet  # noqa: E402


from parity_rules import check_outputs, rel  # noqa: E402


def main():
    """One line per (case, precision): measured errors, flips, how many images had their class checked, and whether the
    rules of tests/parity_rules.py hold (PASS) or which one fails."""
    dev = torch.device("cuda:0")
    only = sys.argv[1:] or list(cases.CASES)
    print(f"{'case':44s} {'mode':5s} {'pooled':>9s} {'logits':>9s} {'map_max':>9s} flips cls_equal cls_checked verdict")
    for name in only:
        case = cases.CASES[name]
        g = dict(np.load(os.path.join(HERE, "golden", name + ".npz")))
        raw = None
        if case["inference"]:
            raw = g["map_sum"] if case["model"] == "count_pipnet" else g["map_max"]
        for precision in ("fp32", "bf16"):
            model = cases.build(case, my_pipnet, my_count)
            sd = {k: v.detach().clone() for k, v in model.state_dict().items()}
            model = model.to(dev)
            model.cpn_precision = precision
            if case["args"].get("activation") == "gumbel_softmax":
                model.cpn_gumbel_noise = cases.gumbel_noise(tuple(int(v) for v in g["map_shape"])).to(dev)
            x = cases.inputs(case).to(dev)
            with torch.no_grad():
                proto, pooled, out = model(x, inference=case["inference"])
            torch.cuda.synchronize()
            pooled, out = pooled.cpu().numpy(), out.cpu().numpy()
            flat = proto.reshape(proto.shape[0], proto.shape[1], -1)
            verdict = "PASS"
            st = dict(flips=-1, class_equal=int((out.argmax(1) == g["pred"]).sum()), class_checked=-1)
            try:
                st = check_outputs(case, sd, precision, pooled, out, g["pooled"], g["out"], raw)
            except AssertionError as exc:
                verdict = "FAIL " + str(exc).replace("\n", " ")[:90]
            print(f"{name:44s} {precision:5s} {rel(pooled, g['pooled']):9.2e} {rel(out, g['out']):9.2e} "
                  f"{rel(flat.amax(2).cpu().numpy(), g['map_max']):9.2e} {st['flips']:5d} "
                  f"{st['class_equal']:5d}/{out.shape[0]:<3d} {st['class_checked']:5d}       {verdict}", flush=True)


if __name__ == "__main__":
    import contextlib
    import io
    # the factories print configuration lines like the reference; keep the table clean
    real = sys.stdout
    buf = io.StringIO()

    class Filter:
        def write(self, s):
            if s.startswith(("Number of", "Using ", "Detected ")) or s == "\n" and getattr(self, "skip", False):
                self.skip = True
                return
            self.skip = False
            real.write(s)

        def flush(self):
            real.flush()

    sys.stdout = Filter()
    main()
