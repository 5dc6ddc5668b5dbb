"""Regenerates docs/API.md from the docstrings of the public Python modules."""
import importlib
import inspect
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
MODULES = ["parallel_kd_tree_b200.ops", "parallel_kd_tree_b200.models", "parallel_kd_tree_b200.parallel.dist_build",
           "parallel_kd_tree_b200.parallel.forest", "parallel_kd_tree_b200.parallel.nccl_baseline",
           "parallel_kd_tree_b200.parallel.bbox", "parallel_kd_tree_b200.utils"]
out = ["# Python API (generated from the docstrings by tools/gen_api_docs.py; the C ABI is documented in "
       "`include/kdtree_b200.h`)\n"]


def doc(o):
    return inspect.cleandoc(o.__doc__) if o.__doc__ else ""


def sig(o):
    try:
        return str(inspect.signature(o))
    except (TypeError, ValueError):
        return "(...)"


for name in MODULES:
    mod = importlib.import_module(name)
    out.append(f"\n## {name}\n")
    if mod.__doc__:
        out.append(doc(mod) + "\n")
    names = getattr(mod, "__all__", None) or [n for n, o in vars(mod).items()
                                              if not n.startswith("_") and getattr(o, "__module__", None) == name]
    for n in names:
        o = getattr(mod, n)
        if inspect.isclass(o):
            out.append(f"### class `{n}{sig(o)}`\n\n{doc(o)}\n")
            for mn, m in vars(o).items():
                if mn.startswith("_") or not (inspect.isfunction(m) or isinstance(m, (classmethod, staticmethod, property))):
                    continue
                f = m.fget if isinstance(m, property) else getattr(m, "__func__", m)
                if f.__doc__:
                    out.append(f"* `{n}.{mn}{'' if isinstance(m, property) else sig(f)}` — {doc(f)}\n")
        elif callable(o):
            out.append(f"### `{n}{sig(o)}`\n\n{doc(o)}\n")
(ROOT / "docs" / "API.md").write_text("\n".join(out))
print("docs/API.md:", sum(len(x.splitlines()) for x in out), "lines")
