"""Reader / writer for the byte format of `flax.serialization.to_bytes` -- the format of the reference's env-model
`.pt` files (/root/reference/train_env_model.py:125-127, /root/reference/utils/envmodel.py:37-51) -- without flax:
a msgpack map whose ndarray leaves are ExtType(1, packb((shape, dtype_name, raw_bytes))) [published flax format;
numpy scalars are ExtType(3), native complex ExtType(2)]. Arrays flax split into chunks (> 2**30 bytes) are rejoined."""
import msgpack
import numpy as np

_EXT_NDARRAY, _EXT_COMPLEX, _EXT_NPSCALAR = 1, 2, 3


def _pack_ndarray(a: np.ndarray) -> bytes:
    a = np.ascontiguousarray(a)
    return msgpack.packb((a.shape, a.dtype.name, a.tobytes("C")), use_bin_type=True)


def _unpack_ndarray(data: bytes) -> np.ndarray:
    shape, dtype_name, buf = msgpack.unpackb(data, raw=True)
    dtype = np.dtype(dtype_name.decode() if isinstance(dtype_name, bytes) else dtype_name)
    return np.frombuffer(buf, dtype=dtype).reshape(tuple(shape)).copy()


def _default(o):
    if isinstance(o, np.ndarray):
        return msgpack.ExtType(_EXT_NDARRAY, _pack_ndarray(o))
    if isinstance(o, np.generic):
        return msgpack.ExtType(_EXT_NPSCALAR, _pack_ndarray(np.asarray(o)))
    if isinstance(o, complex):
        return msgpack.ExtType(_EXT_COMPLEX, msgpack.packb((o.real, o.imag)))
    if hasattr(o, "detach"):  # torch tensor
        return msgpack.ExtType(_EXT_NDARRAY, _pack_ndarray(o.detach().cpu().numpy()))
    raise TypeError(f"cannot serialise {type(o)}")


def _ext_hook(code, data):
    if code == _EXT_NDARRAY:
        return _unpack_ndarray(data)
    if code == _EXT_NPSCALAR:
        return _unpack_ndarray(data)[()]
    if code == _EXT_COMPLEX:
        re, im = msgpack.unpackb(data)
        return complex(re, im)
    return msgpack.ExtType(code, data)


def _unchunk(tree):
    if isinstance(tree, dict):
        if "__msgpack_chunked_array__" in tree:
            shape = tuple(tree["shape"])
            chunks = tree["chunks"]
            flat = np.concatenate([np.ravel(chunks[str(i)]) for i in range(len(chunks))])
            return flat.reshape(shape)
        return {k: _unchunk(v) for k, v in tree.items()}
    return tree


def to_bytes(tree) -> bytes:
    return msgpack.packb(tree, default=_default, strict_types=True)


def from_bytes(data: bytes):
    return _unchunk(msgpack.unpackb(data, ext_hook=_ext_hook, raw=False, strict_map_key=False))
