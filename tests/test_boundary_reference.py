"""The drop-in boundary exercised through the REFERENCE's own seam (build container only: needs /root/reference).

`BaseSystem.get_network` (systems/BaseSystem.py:163-242) looks its classes up on the `raygun.torch.networks` module.  The
reference's BaseSystem.py is executed twice, verbatim, from /root/reference -- once over the reference's networks module,
once with `raygun_b200.networks` installed as `sys.modules["raygun.torch.networks"]` (INTEGRATION.md binding 1) -- for the
network configurations of `default_cycleGAN_conf.json`; the two module trees, their state_dict keys / shapes, and the
initial weights drawn from the same seed must agree.  `raygun_b200.cyclegan.make_generator / make_discriminator` (our own
mirror of get_network, used by bench.py and the GPU tests) must build those same trees: the test fails if it drifts."""
import importlib.util
import json
import os
import re
import sys
import types

import pytest
import torch

from oracle import ref_loader

pytestmark = pytest.mark.skipif(not ref_loader.available(), reason="reference sources not present (GPU box)")

# final_activation: the reference appends whatever it is given to nn.Sequential (BaseSystem.py:221-222), so a NAME
# ("Tanh" -> the class, uninstantiated) raises TypeError there; configs pass an instance ("#torch.nn.Tanh()#").
G_UNET = dict(input_nc=1, output_nc=1, ngf=64, fmap_inc_factor=2, num_downs=3, padding_type="same")
CONFIGS = {
    "unet": ("unet", G_UNET),
    "unet_tanh": ("unet", dict(G_UNET, ngf=16, final_activation=torch.nn.Tanh())),
    "residualunet": ("residualunet", dict(G_UNET, ngf=16)),
    "resnet": ("resnet", dict(input_nc=1, output_nc=1, ngf=16, n_blocks=3, padding_type="same")),
    "patchdiscriminator": ("patchdiscriminator", dict(input_nc=1, ngf=64, n_layers=3, kw=3, downsampling_kw=2)),
}


def _base_system(networks_module, tag):
    """Execute the reference's BaseSystem.py with `raygun.torch.networks` bound to `networks_module`."""
    ref_loader.load()                                   # namespace packages + the reference's own networks
    saved = {k: sys.modules.get(k) for k in ("raygun.torch.networks", "raygun.torch.networks.utils", "raygun.torch.utils",
                                             "raygun.torch.train")}
    rt = sys.modules["raygun.torch"]
    saved_attr = getattr(rt, "networks", None)
    try:
        sys.modules["raygun.torch.networks"] = networks_module
        rt.networks = networks_module
        utils = types.ModuleType("raygun.torch.networks.utils")
        utils.init_weights = networks_module.init_weights
        sys.modules["raygun.torch.networks.utils"] = utils
        tu = types.ModuleType("raygun.torch.utils")
        tu.read_config = lambda *_a, **_k: {}
        sys.modules["raygun.torch.utils"] = tu
        sys.modules["raygun.torch.train"] = types.ModuleType("raygun.torch.train")
        rt.train = sys.modules["raygun.torch.train"]
        path = os.path.join(ref_loader.REF_SRC, "raygun", "torch", "systems", "BaseSystem.py")
        spec = importlib.util.spec_from_file_location(f"_ref_base_system_{tag}", path)
        mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(mod)
        return mod.BaseSystem
    finally:
        for k, v in saved.items():
            if v is None:
                sys.modules.pop(k, None)
            else:
                sys.modules[k] = v
        if saved_attr is not None:
            rt.networks = saved_attr


def _system(cls, ndims=3):
    s = cls.__new__(cls)                                # no config files: get_network only reads self.ndims
    s.ndims = ndims
    return s


def _tree(net):
    # remove_duplicate=False: the reference re-uses ONE padding module instance at several Sequential slots
    return [(name, type(m).__name__.replace("Generator", "")) for name, m in net.named_modules(remove_duplicate=False)]


@pytest.mark.parametrize("name", sorted(CONFIGS))
def test_reference_get_network_builds_the_same_nets_from_our_networks_module(name, capsys):
    import raygun_b200.networks as ours_mod
    from raygun_b200.cyclegan import make_discriminator, make_generator
    ref_ns = ref_loader.load()
    ref_mod = sys.modules["raygun.torch.networks"]
    net_type, kw = CONFIGS[name]
    Ref = _base_system(ref_mod, "ref")
    Ours = _base_system(ours_mod, "ours")
    torch.manual_seed(123)
    ref_net = _system(Ref).get_network(net_type, dict(kw))
    torch.manual_seed(123)
    our_net = _system(Ours).get_network(net_type, dict(kw))
    torch.manual_seed(123)
    mirror = make_discriminator(dict(kw)) if net_type == "patchdiscriminator" else make_generator(net_type, dict(kw))
    inner = our_net[0] if isinstance(our_net, torch.nn.Sequential) else our_net
    assert type(inner).__module__.startswith("raygun_b200."), "the swapped module must supply the classes"
    rsd, osd, msd = ref_net.state_dict(), our_net.state_dict(), mirror.state_dict()
    assert list(rsd) == list(osd) == list(msd)
    for k in rsd:
        assert rsd[k].shape == osd[k].shape == msd[k].shape and rsd[k].dtype == osd[k].dtype, k
        # same RNG stream, same init order -> identical initial weights (kaiming through the reference's init_weights
        # semantics, BaseSystem.py:226-240)
        assert torch.equal(rsd[k], osd[k]), k
        assert torch.equal(rsd[k], msd[k]), k
    # module trees: same names in the same order (class names differ only by our subclasses keeping torch's names)
    rt_, ot_ = _tree(ref_net), _tree(our_net)
    assert [n for n, _ in rt_] == [n for n, _ in ot_]
    for (n, a), (_, b) in zip(rt_, ot_):
        assert a == b or (a, b) in {("ResNet", "ResNet"), ("Sequential", "Generator")} or b.startswith(a) or a.startswith(b), (n, a, b)
    our_net.load_state_dict(rsd)                        # reference checkpoints load strictly
    ref_net.load_state_dict(osd)
    assert [n for n, _ in _tree(mirror)] == [n for n, _ in ot_]


def test_default_cyclegan_conf_network_kwargs_are_accepted():
    """`default_cycleGAN_conf.json` (jsmin-stripped by hand: comments removed) parameterises get_network for the generator
    and the discriminator: the swapped module accepts exactly those kwargs."""
    import raygun_b200.networks as ours_mod
    path = os.path.join(ref_loader.REF_SRC, "raygun", "torch", "default_configs", "default_cycleGAN_conf.json")
    txt = re.sub(r"//[^\n]*", "", open(path).read())
    conf = json.loads(txt)
    Ours = _base_system(ours_mod, "ours_conf")
    s = _system(Ours, conf.get("ndims", 3) or 3)
    gkw = dict(conf["gnet_kwargs"])
    gkw.setdefault("fmap_inc_factor", 2)                # required by UNet.py:359, absent from the default JSON
    if s.ndims == 3:
        g = s.get_network(conf["gnet_type"], gkw)
        d = s.get_network(conf["dnet_type"], dict(conf["dnet_kwargs"]))
        assert isinstance(g, torch.nn.Sequential) and type(g[0]).__module__.startswith("raygun_b200.")
        assert type(d).__module__.startswith("raygun_b200.") or type(d.model[0]).__module__.startswith("raygun_b200.")
