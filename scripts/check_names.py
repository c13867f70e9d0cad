"""Undefined-name check over the Python sources (no linter ships in the image): every name that is loaded must be
bound in an enclosing scope, imported, or a builtin.  Over-approximates bindings, so it reports no false positives on
this tree; it exists for the GPU-only code paths the CPU suite cannot execute.
    python scripts/check_names.py kubeml_b200 bench bench.py baseline/ref_torch.py baseline/reference_arm.py tests"""
import ast, builtins, sys, os
def check(path):
    src = open(path).read()
    tree = ast.parse(src, path)
    problems = []
    builtin_names = set(dir(builtins)) | {"__file__", "__name__", "__doc__", "__package__", "__spec__", "__builtins__"}
    class Scope:
        def __init__(self, parent, kind): self.parent, self.kind, self.names = parent, kind, set()
    def collect_defs(node, scope):
        # names bound in this scope (not descending into nested function/class bodies, but binding their names)
        for child in ast.iter_child_nodes(node):
            bind(child, scope)
    def bind_target(t, scope):
        for n in ast.walk(t):
            if isinstance(n, ast.Name) and isinstance(n.ctx, (ast.Store, ast.Del)):
                scope.names.add(n.id)
    def bind(node, scope):
        if isinstance(node, (ast.FunctionDef, ast.AsyncFunctionDef, ast.ClassDef)):
            scope.names.add(node.name)
            return
        if isinstance(node, ast.Lambda):
            return
        if isinstance(node, (ast.Import, ast.ImportFrom)):
            for a in node.names:
                scope.names.add((a.asname or a.name).split(".")[0])
            return
        if isinstance(node, (ast.Global, ast.Nonlocal)):
            scope.names.update(node.names); return
        if isinstance(node, ast.Name) and isinstance(node.ctx, (ast.Store, ast.Del)):
            scope.names.add(node.id)
        if isinstance(node, ast.ExceptHandler) and node.name:
            scope.names.add(node.name)
        if isinstance(node, (ast.ListComp, ast.SetComp, ast.DictComp, ast.GeneratorExp)):
            # comprehension variables: treat as bound in the enclosing scope (over-approximation)
            for g in node.generators:
                bind_target(g.target, scope)
        if isinstance(node, ast.NamedExpr):
            bind_target(node.target, scope)
        if isinstance(node, ast.MatchAs) and node.name: scope.names.add(node.name)
        for child in ast.iter_child_nodes(node):
            bind(child, scope)
    def visit(node, scope):
        if isinstance(node, (ast.FunctionDef, ast.AsyncFunctionDef, ast.Lambda)):
            s = Scope(scope, "func")
            a = node.args
            for arg in a.posonlyargs + a.args + a.kwonlyargs: s.names.add(arg.arg)
            if a.vararg: s.names.add(a.vararg.arg)
            if a.kwarg: s.names.add(a.kwarg.arg)
            # defaults/decorators evaluated in enclosing scope
            for d in a.defaults + [d for d in a.kw_defaults if d is not None]: visit(d, scope)
            if not isinstance(node, ast.Lambda):
                for d in node.decorator_list: visit(d, scope)
                for st in node.body: bind(st, s)
                for st in node.body: visit(st, s)
            else:
                bind(node.body, s); visit(node.body, s)
            return
        if isinstance(node, ast.ClassDef):
            s = Scope(scope, "class")
            for b in node.bases + [k.value for k in node.keywords] + node.decorator_list: visit(b, scope)
            for st in node.body: bind(st, s)
            for st in node.body: visit(st, s)
            return
        if isinstance(node, ast.Name) and isinstance(node.ctx, ast.Load):
            sc = scope; found = False; first = True
            while sc is not None:
                if node.id in sc.names and (first or sc.kind != "class"):
                    found = True; break
                first = False
                sc = sc.parent
            if not found and node.id not in builtin_names:
                problems.append((node.lineno, node.id))
        for child in ast.iter_child_nodes(node):
            visit(child, scope)
    mod = Scope(None, "module")
    for st in tree.body: bind(st, mod)
    for st in tree.body: visit(st, mod)
    return problems
bad = 0
for root in sys.argv[1:]:
    files = [root] if root.endswith(".py") else [os.path.join(d, f) for d, _, fs in os.walk(root) for f in fs if f.endswith(".py") and "_ref" not in d]
    for f in files:
        for ln, name in check(f):
            print(f"{f}:{ln}: undefined name {name}"); bad += 1
print("problems:", bad)
sys.exit(1 if bad else 0)
