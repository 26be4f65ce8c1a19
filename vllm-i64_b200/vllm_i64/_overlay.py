"""Overlay support: run this package ON TOP OF an installed reference ``vllm_i64``.

``north_star`` keeps the reference's Python module paths so that its engine, scheduler, KV cache, loader, models and
server run unchanged over the new kernels.  This package only holds the hot-path modules; when another ``vllm_i64``
package is importable (the reference: a later ``sys.path`` entry, or the directory named by ``VLLM_I64_REFERENCE``), the
packages here become overlays of it:

  * ``extend(pkg)``  appends the reference's directory of the same (sub)package to ``pkg.__path__`` — modules this repo does
    not provide (``vllm_i64.models``, ``vllm_i64.engine``, ``vllm_i64.core.loader``, ``vllm_i64.layers.rmsnorm`` ...) resolve
    to the reference's files, modules it does provide shadow the reference's — and then executes the reference's own
    ``__init__.py`` in the package namespace, so every re-export the reference defines exists (names this repo binds win);
  * ``fill(module)``  for a module that shadows a reference module of the same name: public names the reference's module
    has and ours lacks (INT4 helpers, flags ...) are taken from the reference's file, loaded under a private name.

Without a reference on the path nothing happens: the package is then the stand-alone hot-path library.
"""
from __future__ import annotations

import importlib.util
import os
import sys
from typing import Optional

_HERE = os.path.dirname(os.path.abspath(__file__))


def reference_root() -> Optional[str]:
    """Directory of the reference's ``vllm_i64`` package, or None."""
    cands = []
    env = os.environ.get("VLLM_I64_REFERENCE")
    if env:
        cands.append(os.path.join(env, "vllm_i64") if not env.rstrip("/").endswith("vllm_i64") else env)
    for p in sys.path:
        if p:
            cands.append(os.path.join(p, "vllm_i64"))
    for c in cands:
        c = os.path.abspath(c)
        if c != _HERE and os.path.isfile(os.path.join(c, "__init__.py")) and os.path.isdir(os.path.join(c, "models")):
            return c
    return None


_REF = reference_root()


def active() -> bool:
    return _REF is not None


def extend(pkg_name: str, pkg_globals: dict) -> None:
    """Overlay package ``pkg_name`` (``vllm_i64`` or ``vllm_i64.<sub>``) on the reference's package of the same name."""
    if _REF is None:
        return
    sub = pkg_name.split(".")[1:]
    ref_dir = os.path.join(_REF, *sub)
    if not os.path.isdir(ref_dir):
        return
    path = pkg_globals["__path__"]
    if ref_dir not in path:
        path.append(ref_dir)
    init = os.path.join(ref_dir, "__init__.py")
    if not os.path.isfile(init):
        return
    ours = {k: v for k, v in pkg_globals.items() if not k.startswith("__")}
    with open(init, "r", encoding="utf-8") as f:
        code = compile(f.read(), init, "exec")
    doc = pkg_globals.get("__doc__")
    exec(code, pkg_globals)                 # the reference's re-exports; its imports resolve through the extended __path__
    pkg_globals.update(ours)                # anything this repo bound in the package namespace wins
    pkg_globals["__doc__"] = doc


def load_reference_module(mod_name: str):
    """The reference's file of module ``mod_name`` (``vllm_i64.a.b``), loaded under ``<mod_name>__reference``; None when
    there is no reference or no such file."""
    if _REF is None:
        return None
    alias = mod_name + "__reference"
    if alias in sys.modules:
        return sys.modules[alias]
    path = os.path.join(_REF, *mod_name.split(".")[1:]) + ".py"
    if not os.path.isfile(path):
        return None
    spec = importlib.util.spec_from_file_location(alias, path)
    mod = importlib.util.module_from_spec(spec)
    sys.modules[alias] = mod
    try:
        spec.loader.exec_module(mod)
    except Exception:
        del sys.modules[alias]
        raise
    return mod


def _guard_int4(name: str, fn):
    """The reference's INT4 functions (group-128 experts, ``int4_linear`` ...) have no sm_100a kernel in this build.  Taken over
    as they are they would run the reference's PyTorch code on the GPU — a fallback nobody asked for.  They stay usable on
    host tensors (load-time packing) and refuse CUDA tensors."""
    import functools

    @functools.wraps(fn)
    def guarded(*args, **kwargs):
        for t in (*args, *kwargs.values()):
            if getattr(t, "is_cuda", False):
                raise NotImplementedError(
                    f"vllm_i64 (B200 build): {name} has no sm_100a kernel; the reference's PyTorch INT4 path is not run in its "
                    "place on CUDA tensors.  Use --quantization int8 or fp8, or no quantisation")
        return fn(*args, **kwargs)
    return guarded


def fill(mod_name: str, mod_globals: dict):
    """Copy the reference module's public names that ``mod_name`` does not define.  Returns the reference module (or None).
    Functions with ``int4`` in their name are wrapped by :func:`_guard_int4`."""
    import types
    ref = load_reference_module(mod_name)
    if ref is None:
        return None
    for k, v in vars(ref).items():
        if k.startswith("__") or k in mod_globals:
            continue
        if isinstance(v, types.FunctionType) and "int4" in k.lower():
            v = _guard_int4(f"{mod_name}.{k}", v)
        mod_globals[k] = v
    return ref


def patch_reference_models() -> list:
    """Classes the reference defines INSIDE its model files (not importable siblings this package could shadow) are
    rebound to the sm_100a modules: ``models.complexity_deep.model.DenseSwiGLUMLP`` (the FFN of dense-306,
    models/complexity_deep/model.py:243-276).  Call once after import, before the model is constructed.  Returns the names
    that were rebound."""
    if _REF is None:
        return []
    done = []
    import importlib
    try:
        m = importlib.import_module("vllm_i64.models.complexity_deep.model")
        from vllm_i64.layers.dense_mlp import DenseSwiGLUMLP
        if getattr(m, "DenseSwiGLUMLP", None) is not DenseSwiGLUMLP:
            m.DenseSwiGLUMLP = DenseSwiGLUMLP
            done.append("vllm_i64.models.complexity_deep.model.DenseSwiGLUMLP")
    except ImportError:
        pass
    done += _hook_reference_loader()
    return done


def _hook_reference_loader() -> list:
    """``vllm-i64 serve --quantization int8`` on a GPU (core/loader.py:627-642): the reference's post-load pass
    ``_quantize_experts`` looks for ``gate_up_proj`` / ``down_proj`` attributes, which ``TokenRoutedMLP`` never had — the
    routed experts stayed in floating point (dead code).  With this package active the pass also switches every
    ``TokenRoutedMLP`` to its W8A8 block (``quantize_w8a8()``), so the flag reaches the INT8 kernels; ``fp8`` does the same
    through ``_quantize_dense_mlp_fp8``.  ``DenseMLP`` needs no hook: the buffers the reference's passes register
    (``gate_up_int8`` …, ``gate_fp8`` …) are picked up by the module as they are."""
    import importlib
    try:
        loader = importlib.import_module("vllm_i64.core.loader")
    except ImportError:
        return []
    if getattr(loader, "_i64b_hooked", False):
        return []
    from vllm_i64.layers.token_routed_mlp import TokenRoutedMLP

    def wrap(name, switch):
        orig = getattr(loader, name, None)
        if orig is None:
            return None

        def hooked(model, *args, **kwargs):
            out = orig(model, *args, **kwargs)
            method = args[0] if args else kwargs.get("method", "fp8")
            for mod in model.modules():
                if isinstance(mod, TokenRoutedMLP):
                    switch(mod, method)
            return out
        hooked.__wrapped__ = orig
        setattr(loader, name, hooked)
        return "vllm_i64.core.loader." + name

    names = [wrap("_quantize_experts", lambda m, method: m.quantize_w8a8() if method == "int8" else None),
             wrap("_quantize_dense_mlp_fp8", lambda m, method: m.quantize_fp8())]
    loader._i64b_hooked = True
    return [n for n in names if n]


def collect_expert_load(model) -> dict:
    """Device-side expert histogram of every token-routed layer's LAST forward, as host lists — what ``/v1/experts`` and
    the ``vllm_i64_experts_load`` gauge report (api/_admin.py:464-562 derive it from ``token_to_expert`` alone today).
    ``{module name: [[count of expert e on route k for e in experts] for k in routes]}``; layers that have not run yet are
    left out.  One small D2H copy per layer: call it from the admin endpoint, not from the step loop."""
    from vllm_i64.layers.token_routed_mlp import TokenRoutedMLP
    out = {}
    for name, mod in model.named_modules():
        if isinstance(mod, TokenRoutedMLP) and getattr(mod, "_block", None) is not None:
            try:
                out[name] = mod.last_expert_counts().cpu().tolist()
            except RuntimeError:
                pass
    return out
