"""Extract the golden traces the reference prints in its demo docstrings.

The reference (mesh-adaptation/movement) stores no numeric fixtures; the only known-answer
vectors for the hot path are the solver traces embedded as comments in ``demos/*.py`` and
the integer/closed-form KATs in ``test/*.py`` (restated directly in tests/test_oracle_golden.py).
This script parses the former into ``tests/golden/demo_traces.json``.  It needs
``/root/reference`` (not present on the GPU box), so its output is committed.

Usage: python tests/golden/make_golden.py [/root/reference]
"""
import json
import os
import re
import sys

REF = sys.argv[1] if len(sys.argv) > 1 else "/root/reference"
HERE = os.path.dirname(os.path.abspath(__file__))

MA_LINE = re.compile(r"#\s+(\d+)\s+Volume ratio\s+([\d.eE+-]+)\s+Variation \(σ/μ\)\s+([\d.eE+-]+)\s+Residual\s+([\d.eE+-]+)")
MV_LINE = re.compile(r"#\s+([\d.]+)(?: s)?\s+Volume ratio\s+([\d.eE+-]+)\s+Variation \(σ/μ\)\s+([\d.eE+-]+)\s+Displacement\s+([\d.eE+-]+) m")
CONV = re.compile(r"#\s+Solver converged in (\d+) iterations?\.")


def parse(path):
    """Return a list of trace blocks; a new block starts whenever the iteration counter resets."""
    blocks, cur, conv = [], None, []
    with open(path, encoding="utf-8") as f:
        for ln, line in enumerate(f, 1):
            m = MA_LINE.match(line)
            if m:
                it = int(m.group(1))
                if cur is None or it == 0:
                    cur = {"kind": "monge_ampere", "first_line": ln, "rows": []}
                    blocks.append(cur)
                cur["rows"].append([it, float(m.group(2)), float(m.group(3)), float(m.group(4))])
                cur["text"] = cur.get("text", []) + [[m.group(2), m.group(3), m.group(4)]]
                continue
            m = MV_LINE.match(line)
            if m:
                t = float(m.group(1))
                if cur is None or cur["kind"] != "forced" or t == 0.0:
                    cur = {"kind": "forced", "first_line": ln, "rows": []}
                    blocks.append(cur)
                cur["rows"].append([t, float(m.group(2)), float(m.group(3)), float(m.group(4))])
                cur["text"] = cur.get("text", []) + [[m.group(2), m.group(3), m.group(4)]]
                continue
            m = CONV.match(line)
            if m and blocks:
                blocks[-1]["converged_in"] = int(m.group(1))
                blocks[-1]["converged_line"] = ln
    return blocks


def main():
    out = {"source": "mesh-adaptation/movement demos (docstring traces)", "files": {}}
    for name in ("monge_ampere_ring.py", "monge_ampere_helmholtz.py", "monge_ampere_periodic.py",
                 "laplacian_smoothing.py", "lineal_spring.py"):
        p = os.path.join(REF, "demos", name)
        out["files"]["demos/" + name] = parse(p)
    with open(os.path.join(HERE, "demo_traces.json"), "w", encoding="utf-8") as f:
        json.dump(out, f, indent=1)
    for k, v in out["files"].items():
        print(k, [(b["kind"], b["first_line"], len(b["rows"]), b.get("converged_in")) for b in v])


if __name__ == "__main__":
    main()
