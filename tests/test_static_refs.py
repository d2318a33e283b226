"""GPU tests, benchmarks and examples cannot run on the CPU box; what can be checked there is
that every ``fm.<name>`` they use exists and every ``from fluxmpi_b200... import name`` resolves,
so a renamed function does not first show up as a failure on GPU time."""
import ast
import importlib
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent


def _files():
    out = list((ROOT / "tests").rglob("*.py")) + list((ROOT / "bench").glob("*.py"))
    out += list((ROOT / "examples").glob("*.py")) + [ROOT / "bench.py", ROOT / "__graft_entry__.py"]
    return sorted(out)


def test_public_names_used_anywhere_exist():
    import fluxmpi_b200 as fm
    missing = []
    for f in _files():
        for n in ast.walk(ast.parse(f.read_text())):
            if isinstance(n, ast.Attribute) and isinstance(n.value, ast.Name) and n.value.id == "fm":
                try:
                    getattr(fm, n.attr)
                except AttributeError:
                    missing.append(f"{f.relative_to(ROOT)}:{n.lineno}: fm.{n.attr}")
    assert not missing, "\n".join(missing)


def test_package_imports_used_anywhere_resolve():
    missing = []
    for f in _files():
        for n in ast.walk(ast.parse(f.read_text())):
            if isinstance(n, ast.ImportFrom) and n.module and n.module.split(".")[0] == "fluxmpi_b200" and n.level == 0:
                try:
                    mod = importlib.import_module(n.module)
                except Exception as e:                                   # noqa: BLE001
                    missing.append(f"{f.relative_to(ROOT)}:{n.lineno}: import {n.module}: {e}")
                    continue
                for a in n.names:
                    if a.name == "*" or hasattr(mod, a.name):
                        continue
                    try:
                        importlib.import_module(f"{n.module}.{a.name}")
                    except Exception:                                    # noqa: BLE001
                        missing.append(f"{f.relative_to(ROOT)}:{n.lineno}: from {n.module} import {a.name}")
    assert not missing, "\n".join(missing)
