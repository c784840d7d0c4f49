"""Step 1 of tools/r_pin/README.md: write the seeded fixture inputs as CSV for an R session."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from ate_replication_causalml_b200 import synth  # noqa: E402

out = sys.argv[1] if len(sys.argv) > 1 else "r_pin_out"
os.makedirs(out, exist_ok=True)
for tag, (n, p0) in {"a": (200, 5), "b": (5000, 21)}.items():
    d = synth.make_dataset(n, p0)
    np.savetxt(os.path.join(out, f"{tag}_x_raw.csv"), d["x_raw"], delimiter=",", fmt="%.17g")
    np.savetxt(os.path.join(out, f"{tag}_w.csv"), d["w"], fmt="%.17g")
    np.savetxt(os.path.join(out, f"{tag}_y.csv"), d["y"], fmt="%.17g")
    np.savetxt(os.path.join(out, f"{tag}_foldid.csv"), synth.fold_ids(n, 10, 7), fmt="%d")
    open(os.path.join(out, f"{tag}_meta.txt"), "w").write(f"n_continuous {d['n_continuous']}\n")
print("wrote inputs to", out)
