"""Generates tests/golden/state_dict_keys.json: the ordered state_dict keys, shapes and dtypes of the REAL reference
models (imported from /root/reference) for every parity case of tests/cases.py -- what a reference checkpoint
(`checkpoint['model_state_dict']`, util/checkpoint_manager.py:66,88,225, loaded with strict=True) contains.
Run in the dev container only:  python tests/golden/make_golden_state_keys.py
"""
import json
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
from make_golden import import_reference  # noqa: E402  (also puts the repo root and tests/ on sys.path)

import cases  # noqa: E402


def main():
    ref_pipnet, ref_count = import_reference()
    table = {}
    for name, case in cases.CASES.items():
        model = cases.build(case, ref_pipnet, ref_count)
        table[name] = [[k, list(v.shape), str(v.dtype)] for k, v in model.state_dict().items()]
    out = os.path.join(HERE, "state_dict_keys.json")
    with open(out, "w") as f:
        json.dump(table, f, indent=0)
    print(out, {k: len(v) for k, v in table.items()})


if __name__ == "__main__":
    main()
