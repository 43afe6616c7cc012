"""Opening the reference's decoder checkpoints with the classes of this package.

The reference training script writes two files per snapshot (main/train_pano2gaussian_decoder.py:297-307):

  * ``decoder{i}.pt``        ``torch.save(decoder.state_dict())`` — keys ``{head}_decoder.backbone.{0,2,4,6}.{weight,bias}``
                             plus the generator's ``G.*`` entries;
  * ``decoder_{i:06d}.pkl``  ``pickle.dump({"training_set_kwargs", "decoder", "dataloader"})`` where the decoder, its five
                             ``Decoder`` heads and the generator are ``persistence.persistent_class`` objects: they pickle
                             as ``torch_utils.persistence._reconstruct_persistent_obj(meta)`` with ``meta`` = class name,
                             the SOURCE TEXT of the defining module, and the instance ``__dict__``
                             (main/torch_utils/persistence.py:37-133, 170-190).

The reference un-pickles by exec-ing that source text, which needs the reference tree (dnnlib, torch_utils, the EG3D /
PanoHead packages) to be importable.  Here nothing from the pickle is executed: persistent objects come back as inert
``PersistentStub`` records, classes that cannot be imported as ``Opaque`` records, and the decoder's constructor arguments
and parameters are then copied into ``ParallelDecoder`` / ``SequentialDecoder`` / ``SequentialDecoderReverse`` of this
package.  The generator is upstream of the path (out of scope): pass your own ``G``; the pickled one is returned as a
stub so that a caller who has the generator code can rebuild it from ``stub.state``.
"""
import io
import pickle

import torch

HEADS = ("xyz", "scale", "rotation", "opacity", "color")


class PersistentStub:
    """What a ``persistent_class`` instance in the pickle turns into: its class name and its ``__dict__``."""

    def __init__(self, class_name, state, module_src=None):
        self.class_name, self.state, self.module_src = class_name, state or {}, module_src

    def __repr__(self):
        return f"<PersistentStub {self.class_name}>"


class Opaque:
    """Stand-in for an object whose class is not importable here (dataloader, config containers, ...)."""

    def __init__(self, *args, **kwargs):
        self.args, self.kwargs, self.state = args, kwargs, None

    def __setstate__(self, state):
        self.state = state

    def __setitem__(self, k, v):          # dict subclasses (dnnlib.EasyDict) are rebuilt item by item
        self.__dict__.setdefault("items_", {})[k] = v

    def __reduce_ex__(self, protocol):
        raise pickle.PicklingError("Opaque placeholders are read-only")


def _reconstruct(meta):
    if meta.get("type") != "class":
        raise pickle.UnpicklingError(f"unsupported persistent object type {meta.get('type')!r}")
    return PersistentStub(meta["class_name"], meta.get("state"), meta.get("module_src"))


class _Unpickler(pickle.Unpickler):
    def find_class(self, module, name):
        if name == "_reconstruct_persistent_obj" and module.split(".")[-1] == "persistence":
            return _reconstruct
        top = module.split(".")[0]
        if top in ("torch", "collections", "numpy", "builtins", "copyreg", "_codecs"):
            return super().find_class(module, name)
        return type(name, (Opaque,), {"__module__": module})


def _module_tensors(obj, prefix=""):
    """state_dict-style ``{dotted name: tensor}`` of a stub / real-module tree (parameters and buffers)."""
    state = obj.state if isinstance(obj, PersistentStub) else obj.__dict__
    out = {}
    for group in ("_parameters", "_buffers"):
        for k, v in (state.get(group) or {}).items():
            if v is not None:
                out[prefix + k] = v.detach() if isinstance(v, torch.Tensor) else v
    for k, child in (state.get("_modules") or {}).items():
        if child is not None:
            out.update(_module_tensors(child, prefix + k + "."))
    return out


def load_state_dict_file(decoder, path_or_dict, map_location="cpu"):
    """Load a reference ``decoder{i}.pt`` (or an already loaded dict) into a decoder of this package.  Every decoder key
    must be present with the right shape; the generator's ``G.*`` entries are returned untouched for the caller."""
    sd = path_or_dict if isinstance(path_or_dict, dict) else torch.load(path_or_dict, map_location=map_location)
    own = {k: v for k, v in sd.items() if not k.startswith("G.")}
    want = {k for k in decoder.state_dict().keys() if not k.startswith("G.")}
    missing, unexpected = sorted(want - set(own)), sorted(set(own) - want)
    if missing or unexpected:
        raise KeyError(f"decoder checkpoint does not match: missing {missing[:4]}, unexpected {unexpected[:4]}")
    decoder.load_state_dict(own, strict=False)
    return {k: v for k, v in sd.items() if k.startswith("G.")}


def load_decoder_pickle(path_or_file, G, device="cuda"):
    """Open a reference ``decoder_*.pkl`` -> (decoder of this package with the checkpoint's parameters, info dict).
    ``info``: ``class_name``, ``init_kwargs``, ``generator`` (PersistentStub of the pickled G or None), ``extra`` (the
    snapshot's other entries, as stubs / opaque records)."""
    from .. import decoder_models
    f = open(path_or_file, "rb") if isinstance(path_or_file, (str, bytes)) else path_or_file
    try:
        snap = _Unpickler(io.BytesIO(f.read())).load()
    finally:
        if isinstance(path_or_file, (str, bytes)):
            f.close()
    stub = snap["decoder"] if isinstance(snap, dict) else snap
    if not isinstance(stub, PersistentStub) or not hasattr(decoder_models, stub.class_name):
        raise TypeError(f"pickle does not hold a known decoder class (got {stub!r})")
    st = stub.state
    kwargs = dict(st.get("_init_kwargs") or {})
    args = list(st.get("_init_args") or ())          # (G, hidden_dim, use_xyz_embedding, use_gen_finetune, device)
    names = ("hidden_dim", "use_xyz_embedding", "use_gen_finetune")
    for i, n in enumerate(names, start=1):
        if n not in kwargs and len(args) > i:
            kwargs[n] = args[i]
    ctor = {n: kwargs[n] for n in names if n in kwargs}
    dec = getattr(decoder_models, stub.class_name)(G, device=device, **ctor)
    tensors = {k: v for k, v in _module_tensors(stub).items() if not k.startswith("G.")}
    load_state_dict_file(dec, tensors)
    gen = (st.get("_modules") or {}).get("G")
    extra = {k: v for k, v in snap.items() if k != "decoder"} if isinstance(snap, dict) else {}
    return dec, dict(class_name=stub.class_name, init_kwargs=ctor, generator=gen if isinstance(gen, PersistentStub) else None,
                     extra=extra)
