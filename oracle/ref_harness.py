"""TEST INFRASTRUCTURE ONLY - imports the UNMODIFIED reference hot path.

In the build container the modules come from ``/root/reference`` (source); on the GPU box, where that
directory does not exist, from ``oracle/_ref`` - the same files byte-compiled by ``oracle/build_ref.py``
(sourceless ``.pyc`` modules, git-ignored, shipped with the snapshot).  Used for: validating the restatement in
``oracle/gn_oracle.py``, generating the golden vectors under ``tests/golden/`` (``oracle/make_golden.py``), the
reference arm of ``bench.py`` and the GPU parity tests that compare against the reference itself.
Nothing in the product package imports this file.

Harness steps:
  1. stub modules for ``addict``, ``plyfile``, ``skimage.measure`` (imported at the top of
     ``reconstruct/utils.py:21-23`` but absent here);
  2. ``Tensor.cuda`` / ``Module.cuda`` -> identity, ``torch.cuda.synchronize/empty_cache`` -> no-op,
     because the reference hard-codes ``.cuda()`` (``optimizer.py:120-136``, ``loss.py:32,63,84``);
  3. ``/root/reference`` and ``/root/reference/reconstruct`` on ``sys.path`` (as ``System.cc:115-116``).
"""
from __future__ import annotations

import os
import sys
import types
import warnings

REFERENCE_ROOT = "/root/reference"
COMPILED_ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "_ref")


def reference_root():
    """Directory to import the reference from: its sources if present, else the byte-compiled copy."""
    if os.path.isfile(os.path.join(REFERENCE_ROOT, "reconstruct", "optimizer.py")):
        return REFERENCE_ROOT
    if os.path.isfile(os.path.join(COMPILED_ROOT, "reconstruct", "optimizer.refbin")):
        return COMPILED_ROOT
    return None


class _RefbinFinder:
    """Meta-path importer for the byte-compiled reference modules of ``oracle/_ref`` (``<pkg>/<mod>.refbin`` = the
    marshalled code object of the unmodified source file; ``<pkg>/__init__.refbin`` makes ``<pkg>`` a package)."""

    def __init__(self, root):
        self.root = root

    def _path(self, fullname):
        base = os.path.join(self.root, *fullname.split("."))
        if os.path.isfile(os.path.join(base, "__init__.refbin")):
            return os.path.join(base, "__init__.refbin"), True
        if os.path.isfile(base + ".refbin"):
            return base + ".refbin", False
        return None, False

    def find_spec(self, fullname, path=None, target=None):
        import importlib.util
        p, is_pkg = self._path(fullname)
        if p is None:
            return None
        spec = importlib.util.spec_from_loader(fullname, self, origin=p, is_package=is_pkg)
        if is_pkg:
            spec.submodule_search_locations = [os.path.dirname(p)]
        return spec

    def create_module(self, spec):
        return None

    def exec_module(self, module):
        import marshal
        with open(module.__spec__.origin, "rb") as f:
            code = marshal.load(f)
        module.__file__ = module.__spec__.origin
        exec(code, module.__dict__)


def available() -> bool:
    return reference_root() is not None


class _AttrDict(dict):
    """Minimal stand-in for ``addict.Dict``: attribute access, nested dicts wrapped."""

    def __init__(self, *a, **kw):
        super().__init__()
        for k, v in dict(*a, **kw).items():
            self[k] = self._wrap(v)

    @classmethod
    def _wrap(cls, v):
        if isinstance(v, dict) and not isinstance(v, cls):
            return cls(v)
        return v

    def __getattr__(self, k):
        try:
            return self[k]
        except KeyError:
            return self.__missing__(k)

    def __missing__(self, k):
        v = type(self)()
        self[k] = v
        return v

    def __setattr__(self, k, v):
        self[k] = self._wrap(v)


_loaded = None
_saved_cuda = None


def patch_cuda():
    """The reference hard-codes ``.cuda()`` (optimizer.py:120-136,151; loss.py:32,63,84,101-109): for the CPU arm
    ``Tensor.cuda`` / ``Module.cuda`` become the identity and ``torch.cuda.synchronize / empty_cache`` no-ops.  These
    are PROCESS-WIDE patches: a process that also drives the GPU (bench.py, the GPU tests) must call
    ``restore_cuda()`` after running the reference - or use ``cpu_only()`` - before it touches CUDA tensors again."""
    global _saved_cuda
    import torch
    if _saved_cuda is None:
        _saved_cuda = (torch.Tensor.cuda, torch.nn.Module.cuda, torch.cuda.synchronize, torch.cuda.empty_cache)
    torch.Tensor.cuda = lambda self, *a, **k: self
    torch.nn.Module.cuda = lambda self, *a, **k: self
    torch.cuda.synchronize = lambda *a, **k: None
    torch.cuda.empty_cache = lambda *a, **k: None


def restore_cuda():
    global _saved_cuda
    import torch
    if _saved_cuda is not None:
        torch.Tensor.cuda, torch.nn.Module.cuda, torch.cuda.synchronize, torch.cuda.empty_cache = _saved_cuda
        _saved_cuda = None


class cpu_only:
    """``with cpu_only(): ...`` - the patches of ``patch_cuda`` for the duration of the block."""

    def __enter__(self):
        patch_cuda()
        return self

    def __exit__(self, *exc):
        restore_cuda()
        return False


def load():
    """Returns a namespace with the reference's modules (imported once)."""
    global _loaded
    if _loaded is not None:
        return _loaded
    root = reference_root()
    if root is None:
        raise RuntimeError("neither /root/reference nor oracle/_ref (python -m oracle.build_ref) is present")
    import torch

    addict = types.ModuleType("addict")
    addict.Dict = _AttrDict
    sys.modules.setdefault("addict", addict)
    sys.modules.setdefault("plyfile", types.ModuleType("plyfile"))
    sk = types.ModuleType("skimage")
    skm = types.ModuleType("skimage.measure")
    sk.measure = skm
    sys.modules.setdefault("skimage", sk)
    sys.modules.setdefault("skimage.measure", skm)

    patch_cuda()

    if root == COMPILED_ROOT:
        sys.meta_path.insert(0, _RefbinFinder(root))
    else:
        for p in (root, os.path.join(root, "reconstruct")):
            if p not in sys.path:
                sys.path.insert(0, p)
    warnings.filterwarnings("ignore")
    import reconstruct.optimizer as r_opt
    import reconstruct.loss as r_loss
    import reconstruct.loss_utils as r_lu
    import reconstruct.utils as r_utils
    import deep_sdf.deep_sdf_decoder as r_dec

    ns = types.SimpleNamespace(optimizer=r_opt, loss=r_loss, loss_utils=r_lu, utils=r_utils,
                               decoder=r_dec, torch=torch, root=root)
    _loaded = ns
    return ns


def build_decoder(state_np):
    """Instantiates the reference ``Decoder`` with the standard DeepSDF specs and loads ``state_np``
    through a ``DataParallel`` wrapper exactly as ``deep_sdf/workspace.py:214-222`` does."""
    ref = load()
    torch = ref.torch
    dec = ref.decoder.Decoder(64, dims=[512] * 8, dropout=list(range(8)), dropout_prob=0.2,
                              norm_layers=list(range(8)), latent_in=[4], weight_norm=True,
                              xyz_in_all=False, use_tanh=False, latent_dropout=False)
    wrapped = torch.nn.DataParallel(dec)
    sd = {"module." + k: torch.from_numpy(v.copy()) for k, v in state_np.items()}
    wrapped.load_state_dict(sd)
    dec = wrapped.module.cpu()          # DataParallel moves the module to cuda:0 on a GPU box; this arm is the CPU path
    dec.eval()
    return dec


def make_configs(cfg_dict):
    ref = load()
    return ref.utils.ForceKeyErrorDict(**cfg_dict)
