"""TEST INFRASTRUCTURE ONLY.  Import the reference's hot-path modules verbatim.

``import raygun`` fails as shipped (gunpowder/daisy absent; dangling
``NotACycle*`` imports).  The hot-path files themselves are
clean PyTorch, so we register empty namespace packages for the ``raygun``
package chain, stub the absent third-party modules, and import the files one
by one in dependency order.

Only usable where ``/root/reference`` exists (the build container).  It is
used to (a) pin ``oracle/restate.py`` and (b) generate ``tests/golden``.
"""
from __future__ import annotations

import importlib
import os
import sys
import types

REF_ROOT = os.environ.get("RAYGUN_REFERENCE", "/root/reference")
REF_SRC = os.path.join(REF_ROOT, "src")


def available() -> bool:
    return os.path.isdir(os.path.join(REF_SRC, "raygun", "torch", "networks"))


def _ns(name: str, path: str | None):
    mod = types.ModuleType(name)
    mod.__path__ = [path] if path else []
    mod.__package__ = name
    sys.modules[name] = mod
    return mod


_loaded = None


def load():
    """Return a namespace with the reference classes (cached)."""
    global _loaded
    if _loaded is not None:
        return _loaded
    if not available():
        raise RuntimeError(f"reference not found under {REF_SRC}")

    rg = os.path.join(REF_SRC, "raygun")
    _ns("raygun", rg)
    _ns("raygun.torch", os.path.join(rg, "torch"))
    for sub in ("networks", "losses", "models", "optimizers"):
        _ns(f"raygun.torch.{sub}", os.path.join(rg, "torch", sub))
    _ns("raygun.evaluation", None)
    va = _ns("raygun.evaluation.validate_affinities", None)
    va.run_validation = lambda *a, **k: None  # only used by MTLSD

    # absent third-party modules that are imported but not used on this path
    for name in ("gunpowder", "daisy", "jsmin"):
        if name not in sys.modules:
            stub = types.ModuleType(name)
            if name == "jsmin":
                stub.jsmin = lambda s: s
            sys.modules[name] = stub
    for name in ("funlib", "funlib.learn", "funlib.learn.torch",
                 "funlib.learn.torch.models", "funlib.learn.torch.models.conv4d"):
        if name not in sys.modules:
            _ns(name, None)
    sys.modules["funlib.learn.torch.models.conv4d"].Conv4d = None

    imp = importlib.import_module
    sys.modules["raygun.utils"] = imp("raygun.utils")

    nets = sys.modules["raygun.torch.networks"]
    nu = imp("raygun.torch.networks.utils")
    for k in ("init_weights", "set_norm_mode", "get_running_norm_stats",
              "get_norm_layers", "NoiseBlock", "ParameterizedNoiseBlock"):
        setattr(nets, k, getattr(nu, k))
    unet = imp("raygun.torch.networks.UNet")
    runet = imp("raygun.torch.networks.ResidualUNet")
    resnet = imp("raygun.torch.networks.ResNet")
    nld = imp("raygun.torch.networks.NLayerDiscriminator")
    nets.UNet, nets.ResidualUNet = unet.UNet, runet.ResidualUNet
    nets.ResNet, nets.NLayerDiscriminator = resnet.ResNet, nld.NLayerDiscriminator

    models = sys.modules["raygun.torch.models"]
    models.BaseModel = imp("raygun.torch.models.BaseModel").BaseModel
    models.FreezableModel = imp("raygun.torch.models.FreezableModel").FreezableModel
    models.CycleModel = imp("raygun.torch.models.CycleModel").CycleModel

    losses = sys.modules["raygun.torch.losses"]
    losses.GANLoss = imp("raygun.torch.losses.GANLoss").GANLoss
    losses.BaseCompetentLoss = imp("raygun.torch.losses.BaseCompetentLoss").BaseCompetentLoss
    losses.LinkCycleLoss = imp("raygun.torch.losses.LinkCycleLoss").LinkCycleLoss
    losses.SplitCycleLoss = imp("raygun.torch.losses.SplitCycleLoss").SplitCycleLoss

    opt = sys.modules["raygun.torch.optimizers"]
    opt.BaseDummyOptimizer = imp("raygun.torch.optimizers.BaseDummyOptimizer").BaseDummyOptimizer

    ns = types.SimpleNamespace(
        UNet=unet.UNet, ConvPass=unet.ConvPass, ConvDownsample=unet.ConvDownsample,
        MaxDownsample=unet.MaxDownsample, Upsample=unet.Upsample,
        ResidualUNet=runet.ResidualUNet, ResNet=resnet.ResNet,
        ResnetGenerator3D=resnet.ResnetGenerator3D, ResnetBlock3D=resnet.ResnetBlock3D,
        NLayerDiscriminator=nld.NLayerDiscriminator,
        NLayerDiscriminator3D=nld.NLayerDiscriminator3D,
        init_weights=nu.init_weights, set_norm_mode=nu.set_norm_mode,
        get_running_norm_stats=nu.get_running_norm_stats,
        NoiseBlock=nu.NoiseBlock, ParameterizedNoiseBlock=nu.ParameterizedNoiseBlock,
        GANLoss=losses.GANLoss, LinkCycleLoss=losses.LinkCycleLoss,
        SplitCycleLoss=losses.SplitCycleLoss, CycleModel=models.CycleModel,
        BaseDummyOptimizer=opt.BaseDummyOptimizer,
    )
    _loaded = ns
    return ns
