"""The product package must never import or execute anything under oracle/ (or transformers):
the oracle is the checker, not a fallback."""
import ast
import os

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG_DIR = os.path.join(ROOT, "distributed-real-time-chat-and-collaboration-tool_b200")
FORBIDDEN = ("oracle", "transformers")


def _imports(path):
    tree = ast.parse(open(path, encoding="utf-8").read(), filename=path)
    for node in ast.walk(tree):
        if isinstance(node, ast.Import):
            for a in node.names:
                yield a.name
        elif isinstance(node, ast.ImportFrom) and node.module and node.level == 0:
            yield node.module


def test_product_python_never_imports_the_oracle_or_transformers():
    offenders = []
    for dirpath, _, files in os.walk(PKG_DIR):
        for f in files:
            if f.endswith(".py"):
                p = os.path.join(dirpath, f)
                for mod in _imports(p):
                    if mod.split(".")[0] in FORBIDDEN:
                        offenders.append((p, mod))
    assert not offenders, offenders


def test_product_native_sources_do_not_reference_the_oracle():
    for dirpath, _, files in os.walk(os.path.join(PKG_DIR, "csrc")):
        for f in files:
            text = open(os.path.join(dirpath, f), encoding="utf-8").read()
            assert "oracle/" not in text and "#include \"../../oracle" not in text, f


def test_missing_library_fails_loudly(tmp_path, monkeypatch):
    import importlib.util

    spec = importlib.util.spec_from_file_location("_lib_copy", os.path.join(PKG_DIR, "_lib.py"))
    mod = importlib.util.module_from_spec(spec)
    monkeypatch.setenv("B200LLM_LIB", str(tmp_path / "libb200llm.so"))  # a path with no library
    src = open(os.path.join(PKG_DIR, "_lib.py"), encoding="utf-8").read()
    try:
        exec(compile(src, "_lib_copy.py", "exec"), mod.__dict__ | {"__file__": os.path.join(PKG_DIR, "_lib.py")})
    except ImportError as exc:
        assert "no CPU fallback" in str(exc).replace("There is no", "no")
    else:
        raise AssertionError("importing without the shared library must raise")
