"""Golden vectors for the part-fit helpers, produced by the REFERENCE's own code in the build container.
    python oracle/gen_golden_parts.py      (needs /root/reference; writes tests/golden/sub_hierarchies.json)
get_sub_hierachies (src/video_mocap/utils/smpl_utils.py:106-160) is executed from its source: the module itself
does not import here (it pulls in moshpp), so the function's source segment is compiled on its own."""
import ast
import itertools
import json
import os
import sys
from typing import List

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import uuo_mocap_b200  # noqa: E402,F401
from uuo_mocap_b200.synthetic import SMPL_PARENTS  # noqa: E402

SRC = "/root/reference/src/video_mocap/utils/smpl_utils.py"
text = open(SRC).read()
fn = [n for n in ast.parse(text).body if isinstance(n, ast.FunctionDef) and n.name == "get_sub_hierachies"][0]
ns = {"np": np, "torch": torch, "itertools": itertools, "List": List}
exec(compile(ast.Module(body=[fn], type_ignores=[]), SRC, "exec"), ns)
out = {"parents": [int(p) for p in SMPL_PARENTS], "subtrees": {}}
for k in range(1, 7):
    out["subtrees"][str(k)] = ns["get_sub_hierachies"](np.array(SMPL_PARENTS), k)
json.dump(out, open(os.path.join(ROOT, "tests", "golden", "sub_hierarchies.json"), "w"))
print({k: len(v) for k, v in out["subtrees"].items()})
