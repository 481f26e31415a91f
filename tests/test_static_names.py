"""Static check that no function in the repository loads a name that is bound nowhere it can see (builtins, module
level, its own or an enclosing function's scope).  Code paths that only run on a GPU box -- bench.py's reference arm,
the tools -- are never executed by the CPU suite; a stray name there (it happened: a key copied from one JSON payload
into the other arm's) would otherwise first fail on the driver's box."""
import ast
import builtins
import glob
import os

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _bound_in(fn):
    b = set()
    for n in ast.walk(fn):
        if isinstance(n, ast.Name) and isinstance(n.ctx, (ast.Store, ast.Del)):
            b.add(n.id)
        elif isinstance(n, ast.arg):
            b.add(n.arg)
        elif isinstance(n, (ast.Import, ast.ImportFrom)):
            b.update((a.asname or a.name).split(".")[0] for a in n.names)
        elif isinstance(n, (ast.FunctionDef, ast.AsyncFunctionDef, ast.ClassDef)):
            b.add(n.name)
        elif isinstance(n, ast.ExceptHandler) and n.name:
            b.add(n.name)
    return b


def _module_level(tree):
    b = set()

    def walk(node):
        for ch in ast.iter_child_nodes(node):
            if isinstance(ch, (ast.FunctionDef, ast.AsyncFunctionDef, ast.ClassDef)):
                b.add(ch.name)
                continue                      # names bound inside a function are not module names
            if isinstance(ch, ast.Lambda):
                continue
            if isinstance(ch, ast.Name) and isinstance(ch.ctx, (ast.Store, ast.Del)):
                b.add(ch.id)
            elif isinstance(ch, (ast.Import, ast.ImportFrom)):
                b.update((a.asname or a.name).split(".")[0] for a in ch.names)
            elif isinstance(ch, ast.ExceptHandler) and ch.name:
                b.add(ch.name)
            walk(ch)

    walk(tree)
    return b


def unbound_names(path):
    tree = ast.parse(open(path).read(), filename=path)
    visible = set(dir(builtins)) | _module_level(tree) | {"__class__", "__file__", "__name__", "__doc__"}
    found = []

    def rec(node, outer):
        for ch in ast.iter_child_nodes(node):
            if isinstance(ch, (ast.FunctionDef, ast.AsyncFunctionDef, ast.Lambda)):
                scope = outer | _bound_in(ch)
                found.extend((path, n.lineno, n.id) for n in ast.walk(ch)
                             if isinstance(n, ast.Name) and isinstance(n.ctx, ast.Load) and n.id not in scope)
                rec(ch, scope)
            else:
                rec(ch, outer)

    rec(tree, visible)
    return sorted(set(found))


def test_no_unbound_names_anywhere():
    files = []
    for pat in ("*.py", "distributed_pytorch_from_scratch_b200/**/*.py", "tools/*.py", "tests/*.py", "models/*.py"):
        files += glob.glob(os.path.join(REPO, pat), recursive=True)
    assert len(files) > 60
    bad = [hit for f in files for hit in unbound_names(f)]
    assert not bad, bad


def test_checker_catches_a_stray_name(tmp_path):
    p = tmp_path / "m.py"
    p.write_text("def a():\n    parity = 1\n    return parity\n\n\ndef b():\n    return {'parity': parity}\n")
    assert [(ln, name) for _, ln, name in unbound_names(str(p))] == [(7, "parity")]
