"""persistence - load the reference's on-disk formats into this package's modules.

Formats (all produced by the reference, none of its code is needed or executed to read them):

  * StyleGAN2-ADA network pickles - ``pickle.dump(dict(G=..., D=..., G_ema=...))`` (training_loop.py:361-371) and
    ``torch.save(G_ema)`` files such as VisCoIN's ``gan-cub.pkl`` (viscoin/cli/train.py:204).  Every network object in
    them is a *persistent class* instance (torch_utils/persistence.py:35-130): it pickles as
    ``_reconstruct_persistent_obj(meta)`` with ``meta = {type, version, module_src, class_name, state}`` and the
    reference rebuilds it by ``exec``-ing the pickled ``module_src`` (persistence.py:179-227).  Here the source is
    NEVER executed: ``class_name`` is looked up in a table of this package's own classes, the object is constructed
    from the recorded ``_init_args / _init_kwargs`` and the recorded parameters / buffers are copied in.  Pre-trained
    ``--cfg=auto`` networks carry ``num_fp16_res=4, conv_clamp=256`` (stylegan2_ada/train.py:181-182) in those kwargs:
    fp16 blocks map onto the bf16-resident pipeline with the clamp applied in the epilogue (networks.SynthesisBlock).
  * whole-object ``torch.save`` pickles of the VisCoIN models (viscoin/models/utils.py:211-233: ``VisCoINModels`` with
    ``GeneratorAdapted``, ``ConceptExtractor``, ``Explainer``, ``Classifier``) - plain (non-persistent) classes pickled by
    qualified name: mapped onto this package's classes by name, rebuilt from their recorded attributes.
  * VisCoIN ``.pth`` checkpoints (viscoin/models/utils.py:177-208): a dict of four ``state_dict``s -
    ``load_viscoin_checkpoint``.

The unpickler is RESTRICTED: globals resolve only from an allow-list (torch / numpy / collections / transformers
internals that tensors and library modules need) or from the name tables below; anything else raises.  Classes that are
not on the synthesis path (Discriminator & co.) come back as inert ``OpaqueObject`` records.
"""

import collections
import io
import pickle
import zipfile

import torch

from . import networks as _networks
from . import viscoin_models as _models


class EasyDict(dict):
    """dnnlib.util.EasyDict (dnnlib/util.py:40-52): attribute access to dict items."""

    def __getattr__(self, name):
        try:
            return self[name]
        except KeyError:
            raise AttributeError(name)

    def __setattr__(self, name, value):
        self[name] = value

    def __delattr__(self, name):
        del self[name]


class OpaqueObject:
    """A pickled object of a class this package has no counterpart for (Discriminator, AugmentPipe, ...): its class name
    and raw state, nothing executable."""

    def __init__(self, class_name, state):
        self.class_name, self.state = class_name, state

    def __repr__(self):
        return "OpaqueObject(%s)" % self.class_name


# persistent classes of stylegan2_ada/training/networks.py -> this package
PERSISTENT_CLASSES = {
    "FullyConnectedLayer": _networks.FullyConnectedLayer, "Conv2dLayer": _networks.Conv2dLayer,
    "MappingNetwork": _networks.MappingNetwork, "SynthesisLayer": _networks.SynthesisLayer,
    "ToRGBLayer": _networks.ToRGBLayer, "SynthesisBlock": _networks.SynthesisBlock,
    "SynthesisNetwork": _networks.SynthesisNetwork, "Generator": _networks.Generator,
}


def _module_state(state):
    """(parameters, buffers, children, training) of a pickled torch.nn.Module __dict__."""
    return (state.get("_parameters") or {}, state.get("_buffers") or {}, state.get("_modules") or {}, state.get("training", True))


def _fill(obj, state, what):
    """Copy recorded parameters / buffers into a freshly constructed module and adopt the (already rebuilt) children."""
    params, buffers, children, training = _module_state(state)
    with torch.no_grad():
        for name, value in params.items():
            if value is None:
                continue
            tgt = obj._parameters.get(name)
            if tgt is None or tuple(tgt.shape) != tuple(value.shape):
                raise pickle.UnpicklingError("%s: parameter %r of shape %s does not fit %s" % (
                    what, name, tuple(value.shape), None if tgt is None else tuple(tgt.shape)))
            tgt.copy_(value.detach().to(tgt.dtype))
            tgt.requires_grad_(bool(getattr(value, "requires_grad", True)))
        for name, value in buffers.items():
            if value is None:
                continue
            if name not in obj._buffers:
                raise pickle.UnpicklingError("%s: unexpected buffer %r" % (what, name))
            obj._buffers[name] = value.detach().clone().to(obj._buffers[name].dtype) if obj._buffers[name] is not None else value.detach().clone()
    for name, child in children.items():
        if isinstance(child, OpaqueObject):
            raise pickle.UnpicklingError("%s: child %r (%s) has no counterpart in this package" % (what, name, child.class_name))
        if name not in obj._modules:
            raise pickle.UnpicklingError("%s: unexpected child module %r" % (what, name))
        obj._modules[name] = child
    obj.train(bool(training))
    return obj


def _reconstruct_persistent_obj(meta):
    """Stand-in for torch_utils.persistence._reconstruct_persistent_obj (persistence.py:179-202) that never executes
    ``meta['module_src']``."""
    if meta.get("type") != "class":
        raise pickle.UnpicklingError("persistent object of type %r" % (meta.get("type"),))
    name, state = meta["class_name"], meta["state"] or {}
    cls = PERSISTENT_CLASSES.get(name)
    if cls is None:
        return OpaqueObject(name, state)
    args, kwargs = tuple(state.get("_init_args", ())), dict(state.get("_init_kwargs", {}))
    for k in ("mapping_kwargs", "synthesis_kwargs"):
        if k in kwargs:
            kwargs[k] = dict(kwargs[k])
    try:
        obj = cls(*args, **kwargs)
    except (AssertionError, TypeError) as e:
        # a configuration this package does not implement (e.g. the Discriminator's Conv2dLayer(down=2)): keep it inert.  If
        # the synthesis path needs it, adopting it as a child fails loudly in _fill.
        return OpaqueObject(name, dict(state, _unsupported=str(e)))
    obj._init_args, obj._init_kwargs = args, kwargs
    return _fill(obj, state, name)


class _Shim(torch.nn.Module):
    """Unpickling target for the reference's plain (non-persistent) model classes: receives the pickled ``__dict__`` and is
    converted into this package's class by ``_convert`` once the whole file is loaded."""
    _target = None

    def __setstate__(self, state):
        self.__dict__["_shim_state"] = state


def _shim(name):
    return type("_Shim_" + name, (_Shim,), {"_target": name})


class VisCoINModels:
    """viscoin/models/utils.py:20-27 (the dataclass of the four models)."""

    def __init__(self, classifier=None, concept_extractor=None, explainer=None, gan=None):
        self.classifier, self.concept_extractor, self.explainer, self.gan = classifier, concept_extractor, explainer, gan

    def __setstate__(self, state):
        self.__dict__.update(state)


def _convert(obj):
    """Replace shims (recursively through containers / bundles) by this package's modules."""
    if isinstance(obj, _Shim):
        st = obj.__dict__["_shim_state"]
        params, buffers, children, training = _module_state(st)
        children = {k: _convert(v) for k, v in children.items()}
        st = dict(st, _modules=children)
        t = obj._target
        if t == "GeneratorAdapted":
            syn = children["synthesis"]
            mp = children["mapping"]
            new = _networks.GeneratorAdapted.__new__(_networks.GeneratorAdapted)
            torch.nn.Module.__init__(new)
            new.z_dim, new.w_dim = st.get("z_dim", 256), st.get("w_dim", 512)
            new.img_resolution, new.img_channels = syn.img_resolution, syn.img_channels
            new.synthesis, new.mapping, new.num_ws = syn, mp, syn.num_ws
            new.train(bool(training))
            return new
        if t == "MappingNetworkAdapted":
            first = children["s0"]
            layers = list(first._modules.values()) if isinstance(first, torch.nn.ModuleList) else list(first)
            new = _networks.MappingNetworkAdapted(st["z1_dim"], st["z2_dim"], st["w_dim"], st["num_ws"], num_layers=st.get("num_layers", len(layers)),
                                                  coarse_layer=st.get("coarse_layer", 2), mid_layer=st.get("mid_layer", 10))
            return _fill(new, st, t)
        if t == "ConceptExtractor":
            c1, c5 = children["conv1"], children["conv5"]
            new = _models.ConceptExtractor(n_concepts=c5.out_channels, latent_channels=c1.out_channels, input_channels1=c1.in_channels,
                                           input_channels2=children["conv2"].in_channels, input_channels3=children["conv3"].in_channels)
            return _fill(new, st, t)
        if t == "Explainer":
            lin = children["linear"]
            new = _models.Explainer(n_concepts=lin.in_features, n_classes=lin.out_features)
            return _fill(new, st, t)
        if t == "Classifier":
            new = _models.Classifier.__new__(_models.Classifier)
            torch.nn.Module.__init__(new)
            new.resnet, new.linear = children["resnet"], children["linear"]
            new.last_size = st.get("last_size", new.linear.in_features)
            new.train(bool(training))
            return new
        raise pickle.UnpicklingError("no conversion for %s" % t)
    if isinstance(obj, VisCoINModels):
        for k, v in list(obj.__dict__.items()):
            setattr(obj, k, _convert(v))
        return obj
    if isinstance(obj, torch.nn.ModuleList):
        for k, v in list(obj._modules.items()):
            obj._modules[k] = _convert(v)
        return obj
    if isinstance(obj, dict):
        for k in list(obj.keys()):
            obj[k] = _convert(obj[k])
        return obj
    if isinstance(obj, (list, tuple)):
        return type(obj)(_convert(v) for v in obj)
    return obj


_SHIMS = {n: _shim(n) for n in ("GeneratorAdapted", "MappingNetworkAdapted", "ConceptExtractor", "Explainer", "Classifier")}

# qualified names the reference's pickles use -> what they resolve to here
NAME_TABLE = {
    ("torch_utils.persistence", "_reconstruct_persistent_obj"): _reconstruct_persistent_obj,
    ("stylegan2_ada.torch_utils.persistence", "_reconstruct_persistent_obj"): _reconstruct_persistent_obj,
    ("dnnlib.util", "EasyDict"): EasyDict,
    ("stylegan2_ada.dnnlib.util", "EasyDict"): EasyDict,
    ("viscoin.models.gan", "GeneratorAdapted"): _SHIMS["GeneratorAdapted"],
    ("viscoin.models.gan", "MappingNetworkAdapted"): _SHIMS["MappingNetworkAdapted"],
    ("viscoin.models.concept_extractors", "ConceptExtractor"): _SHIMS["ConceptExtractor"],
    ("viscoin.models.explainers", "Explainer"): _SHIMS["Explainer"],
    ("viscoin.models.classifiers", "Classifier"): _SHIMS["Classifier"],
    ("viscoin.models.utils", "VisCoINModels"): VisCoINModels,
}
_ALLOWED_PREFIXES = ("torch.", "numpy.", "transformers.", "torchvision.transforms.")
_ALLOWED_EXACT = {("collections", "OrderedDict"), ("collections", "defaultdict"), ("builtins", "set"), ("builtins", "frozenset"),
                  ("builtins", "slice"), ("builtins", "complex"), ("builtins", "getattr"), ("copyreg", "_reconstructor"),
                  ("builtins", "object"), ("numpy", "dtype"), ("numpy", "ndarray"), ("torch", "Size"), ("torch", "device"),
                  ("_codecs", "encode")}


class RestrictedUnpickler(pickle.Unpickler):
    def find_class(self, module, name):
        if module == "__builtin__":  # protocol-2 streams (torch.save's default) name Python-2 style modules
            module = "builtins"
        hit = NAME_TABLE.get((module, name))
        if hit is not None:
            return hit
        if (module, name) in _ALLOWED_EXACT or module == "torch" or module.startswith(_ALLOWED_PREFIXES):
            if name.startswith("__") or "system" in name or name in ("eval", "exec", "open", "load", "loads"):
                raise pickle.UnpicklingError("refused global %s.%s" % (module, name))
            return super().find_class(module, name)
        if module.endswith("training.networks") and name in PERSISTENT_CLASSES:
            return PERSISTENT_CLASSES[name]
        raise pickle.UnpicklingError("global %s.%s is not on the allow-list of viscoin_next_b200.persistence" % (module, name))


class _PickleModule:
    """The `pickle_module` torch.load() drives: same entry points as `pickle`, restricted Unpickler."""
    __name__ = "viscoin_next_b200.persistence"
    Unpickler = RestrictedUnpickler
    Pickler = pickle.Pickler
    UnpicklingError = pickle.UnpicklingError

    @staticmethod
    def load(f, **kw):
        return RestrictedUnpickler(f, **kw).load()

    @staticmethod
    def loads(b, **kw):
        return RestrictedUnpickler(io.BytesIO(b), **kw).load()


def load_reference_pickle(path_or_file, map_location="cpu"):
    """Load a reference pickle (see the module docstring) and return what it holds with every network object replaced by
    this package's module (``Generator`` / ``SynthesisNetwork`` / ``GeneratorAdapted`` / ...), parameters and buffers as
    recorded.  ``dict(G=..., D=..., G_ema=...)`` snapshots come back as a dict whose D / augment_pipe are OpaqueObjects."""
    if isinstance(path_or_file, (str, bytes)) or hasattr(path_or_file, "__fspath__"):
        with open(path_or_file, "rb") as fh:
            data = fh.read()
    else:
        data = path_or_file.read()
    if zipfile.is_zipfile(io.BytesIO(data)):  # torch.save (zip container): storages are separate records
        obj = torch.load(io.BytesIO(data), map_location=map_location, pickle_module=_PickleModule, weights_only=False)
    else:
        obj = RestrictedUnpickler(io.BytesIO(data)).load()
        if isinstance(obj, int):  # torch.save's legacy (pre-zip) stream begins with its magic number
            obj = torch.load(io.BytesIO(data), map_location=map_location, pickle_module=_PickleModule, weights_only=False)
    obj = _convert(obj)
    return _to(obj, map_location) if map_location is not None else obj


def _to(obj, device):
    """Rebuilt modules are constructed on the CPU (their recorded tensors may already be on `device`): move everything."""
    if isinstance(obj, torch.nn.Module) or torch.is_tensor(obj):
        return obj.to(device)
    if isinstance(obj, VisCoINModels):
        for k, v in list(obj.__dict__.items()):
            setattr(obj, k, _to(v, device))
    elif isinstance(obj, dict):
        for k in list(obj.keys()):
            obj[k] = _to(obj[k], device)
    return obj


def load_generator(path_or_file, key="G_ema", map_location="cpu"):
    """The synthesis-carrying generator of a reference pickle: the object itself (``gan-cub.pkl``) or entry `key` of a
    training snapshot."""
    obj = load_reference_pickle(path_or_file, map_location=map_location)
    if isinstance(obj, dict):
        obj = obj[key]
    if not isinstance(obj, _networks.Generator):
        raise TypeError("expected a Generator, found %r" % type(obj).__name__)
    return obj


def load_viscoin_checkpoint(path_or_file, classifier, concept_extractor, explainer, gan, map_location="cpu"):
    """viscoin/models/utils.py:195-208: a dict of the four state_dicts, loaded strictly."""
    ck = torch.load(path_or_file, map_location=map_location, weights_only=True)
    classifier.load_state_dict(ck["classifier"])
    concept_extractor.load_state_dict(ck["concept_extractor"])
    explainer.load_state_dict(ck["explainer"])
    gan.load_state_dict(ck["gan"])
    return ck.keys()
