"""Function registry: what `kubeml function create/delete/list` manage.

The reference turns a single Python file into a Fission Package + Function + HTTP trigger
(ml/pkg/kubeml-cli/cmd/function.go:71-145) and the environment container later "specialises" a
warm pod by importing that file and grabbing `main` (ml/environment/server.py:60-110).  Here a
function is the same single source file, copied under `$KUBEML_HOME/functions/<name>.py`;
specialising is importing it into the worker process.  Built-in example functions
(kubeml_b200.functions.*) are addressable as `builtin:<module>` or by bare module name.
"""
from __future__ import annotations

import importlib
import importlib.util
import os
import shutil
import sys
import threading
from typing import Callable, Dict, List

from ..api.const import home
from ..exceptions import KubeMLException

_BUILTIN_PKG = "kubeml_b200.functions"


class FunctionRegistry:
    def __init__(self, root: str | None = None):
        self.root = root or os.path.join(home(), "functions")
        os.makedirs(self.root, exist_ok=True)
        self._cache: Dict[str, tuple] = {}
        self._lock = threading.Lock()

    def _path(self, name: str) -> str:
        if not name or "/" in name or name.startswith("."):
            raise KubeMLException(f"invalid function name {name!r}", 400)
        return os.path.join(self.root, f"{name}.py")

    def create(self, name: str, code_path: str) -> None:
        if not os.path.isfile(code_path):
            raise KubeMLException(f"code file {code_path} does not exist", 400)
        if os.path.exists(self._path(name)):
            raise KubeMLException(f"function {name} already exists", 400)
        shutil.copyfile(code_path, self._path(name))

    def delete(self, name: str) -> None:
        try:
            os.remove(self._path(name))
        except FileNotFoundError:
            raise KubeMLException(f"function {name} does not exist", 404)
        with self._lock:
            self._cache.pop(name, None)

    def list(self) -> List[str]:
        return sorted(f[:-3] for f in os.listdir(self.root) if f.endswith(".py"))

    def builtins(self) -> List[str]:
        import pkgutil
        pkg = importlib.import_module(_BUILTIN_PKG)
        return sorted(m.name for m in pkgutil.iter_modules(pkg.__path__)
                      if m.name.startswith("function_"))

    def exists(self, name: str) -> bool:
        if os.path.isfile(self._path(name.split(":", 1)[-1])) and not name.startswith("builtin:"):
            return True
        return self._builtin_module(name) is not None

    def _builtin_module(self, name: str):
        short = name.split(":", 1)[1] if name.startswith("builtin:") else name
        for cand in (short, f"function_{short}"):
            try:
                if importlib.util.find_spec(f"{_BUILTIN_PKG}.{cand}") is not None:
                    return f"{_BUILTIN_PKG}.{cand}"
            except (ImportError, ValueError):
                continue
        return None

    def load(self, name: str, entrypoint: str = "main") -> Callable:
        """Specialise: return the function's entrypoint, re-importing when the file changed."""
        if not name.startswith("builtin:") and os.path.isfile(self._path(name)):
            path = self._path(name)
            mtime = os.path.getmtime(path)
            with self._lock:
                hit = self._cache.get(name)
                if hit and hit[0] == mtime:
                    return getattr(hit[1], entrypoint)
                spec = importlib.util.spec_from_file_location(f"kubeml_userfunc_{name}", path)
                mod = importlib.util.module_from_spec(spec)
                sys.modules[spec.name] = mod
                try:
                    spec.loader.exec_module(mod)
                except Exception as e:  # noqa: BLE001
                    raise KubeMLException(f"could not load function {name}: {e!r}", 500)
                self._cache[name] = (mtime, mod)
                if not hasattr(mod, entrypoint):
                    raise KubeMLException(f"function {name} has no {entrypoint}()", 500)
                return getattr(mod, entrypoint)
        modname = self._builtin_module(name)
        if modname is None:
            raise KubeMLException(f"function {name} does not exist", 404)
        mod = importlib.import_module(modname)
        return getattr(mod, entrypoint)
