"""`generate_cond.py` / `generate_steps.py` of the reference against the drop-in modules, statically
(build container only: the reference tree does not travel to the GPU box; the same call
sequences are EXECUTED on the B200 by tests/test_gpu_scripts.py).

With the product directory ahead of /root/reference on sys.path -- how a user drops the
replacement in -- the reference's scripts must bind every name they import to the product's
classes, and every call they make must be accepted by the product's signatures
(generate_cond.py:41-77, generate_steps.py:97-189)."""
import ast
import importlib.util
import inspect
import os
import sys

import pytest

from oracle import load_reference

REF = load_reference.REFERENCE_ROOT
pytestmark = pytest.mark.skipif(not load_reference.available(), reason="reference tree not present")


def _product():
    import diff
    import entityCsvSampler
    import utils
    from models import unet_cond, unet_cond_geom, vae
    return {"Diffuser": diff.Diffuser, "EntityCsvSampler": entityCsvSampler.EntityCsvSampler, "Utils": utils.Utils,
            "UnetCond": unet_cond.UnetCond, "UnetCondWithGeomHead": unet_cond_geom.UnetCondWithGeomHead,
            "VAE": vae.VAE}


def test_generate_steps_imports_bind_to_the_product():
    """generate_steps.py has a __main__ guard, so the unmodified file can be imported: with the
    product first on sys.path its `from diff import Diffuser` etc. must yield the product."""
    prod = _product()
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    assert sys.path.index(os.path.join(root, "diffusion-model_b200")) < (
        sys.path.index(REF) if REF in sys.path else len(sys.path))
    sys.path.append(REF)          # second, as a user would have it
    try:
        spec = importlib.util.spec_from_file_location("_ref_generate_steps", os.path.join(REF, "generate_steps.py"))
        gs = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(gs)
    finally:
        sys.path.remove(REF)
    for name in ("Diffuser", "EntityCsvSampler", "UnetCondWithGeomHead", "VAE", "Utils"):
        assert getattr(gs, name) is prod[name], name
        assert getattr(gs, name).__module__ in ("diff", "entityCsvSampler", "utils", "models.unet_cond_geom",
                                                "models.vae")
    # the function the script calls exists with the reference's keyword-only signature
    sig = inspect.signature(gs.save_reverse_steps_for_csv_row)
    assert {"csv_path", "row_index", "class_id", "model", "vae", "device", "num_timesteps", "z_shape",
            "guidance_scale", "null_label", "save_steps", "save_every"} <= set(sig.parameters)


def _calls(path):
    tree = ast.parse(open(path, encoding="utf-8").read())
    out = []
    for node in ast.walk(tree):
        if isinstance(node, ast.Call):
            out.append((ast.unparse(node.func), len(node.args), [k.arg for k in node.keywords], node.lineno))
    return tree, out


def _imports(tree):
    names = {}
    for node in ast.walk(tree):
        if isinstance(node, ast.ImportFrom):
            for a in node.names:
                names[a.asname or a.name] = node.module
    return names


def _accepts(fn, npos, kws, bound_self=False):
    sig = inspect.signature(fn)
    args = [object()] * (npos + (1 if bound_self else 0))
    sig.bind(*args, **{k: object() for k in kws})     # raises TypeError if the call would


@pytest.mark.parametrize("script", ["generate_cond.py", "generate_steps.py"])
def test_every_call_site_is_accepted_by_the_product_signatures(script):
    prod = _product()
    tree, calls = _calls(os.path.join(REF, script))
    imported = _imports(tree)
    for name, module in imported.items():
        if module in ("diff", "entityCsvSampler", "utils", "models.unet_cond", "models.unet_cond_geom", "models.vae"):
            assert name in prod, (script, name, module)
    # variable -> class, as the scripts assign them
    D, S, U = prod["Diffuser"], prod["EntityCsvSampler"], prod["Utils"]
    table = {
        "Diffuser": (D.__init__, True), "EntityCsvSampler": (S.__init__, True),
        "UnetCondWithGeomHead": (prod["UnetCondWithGeomHead"].__init__, True), "VAE": (prod["VAE"].__init__, True),
        "Utils.loadModel": (U.loadModel, False), "Utils.saveImages": (U.saveImages, False),
        "sampler.sample": (S.sample, True), "sampler.set_class_id": (S.set_class_id, True),
        "sampler.load_cond": (S.load_cond, True),
        "diffuser.denoise_cond": (D.denoise_cond, True), "diffuser.reverse_to_img": (D.reverse_to_img, True),
        "vae.decode": (prod["VAE"].decode, True), "vae.eval": (prod["VAE"].eval, True),
        "model.to": (prod["UnetCondWithGeomHead"].to, True), "model_noise.eval": (prod["UnetCondWithGeomHead"].eval, True),
    }
    seen = set()
    for func, npos, kws, line in calls:
        if func in table:
            fn, bound = table[func]
            try:
                _accepts(fn, npos, kws, bound_self=bound)
            except TypeError as e:
                raise AssertionError(f"{script}:{line}: {func}({npos} positional, {kws}) rejected: {e}")
            seen.add(func)
    want = {"generate_cond.py": {"Diffuser", "EntityCsvSampler", "UnetCondWithGeomHead", "VAE", "Utils.loadModel",
                                 "Utils.saveImages", "sampler.sample", "sampler.set_class_id"},
            "generate_steps.py": {"Diffuser", "EntityCsvSampler", "sampler.load_cond", "diffuser.denoise_cond",
                                  "diffuser.reverse_to_img", "vae.decode", "vae.eval", "model.to", "model_noise.eval",
                                  "Utils.loadModel", "UnetCondWithGeomHead", "VAE"}}[script]
    assert want <= seen, want - seen


def test_attributes_the_scripts_read_exist():
    """Diffuser.device / num_timesteps / betas / alphas / alpha_bars (diff.py:11-16); the string
    device the scripts pass is accepted at construction on a machine without a GPU too."""
    prod = _product()
    d = prod["Diffuser"](num_timesteps=1000, device="cpu")
    for a in ("num_timesteps", "device", "betas", "alphas", "alpha_bars"):
        assert hasattr(d, a)
    s = prod["EntityCsvSampler"](diffuser=d, model=None, vae=None, class_id=1)
    assert s.class_id == 1 and s.base_wh == (400, 400)
    s.set_class_id(3)
    assert s.class_id == 3
