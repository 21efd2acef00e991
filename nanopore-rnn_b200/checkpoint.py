"""Read a TensorFlow checkpoint "bundle" (``<prefix>.index`` + ``<prefix>.data-00000-of-00001``)
without TensorFlow, and normalise variable names to the ones the forward pass looks up.

Replaces ``tf.train.Saver.restore`` at run_nanotensor.py:337-338 / :367-368 of the reference.
Format (tensorflow/core/util/tensor_bundle + leveldb table format, restated from the published
format; verified on the two ``.index`` files the reference ships):

* ``.index`` is an SSTable: 48-byte footer = varint64 metaindex(offset,size), index(offset,size),
  padding, 8-byte magic.  Each block is followed by a 1-byte compression type (0 = none) and a
  4-byte CRC.  A block holds prefix-compressed entries (varint shared, non_shared, value_len,
  key suffix, value) and ends with a uint32 restart array + uint32 count.
* key ``""`` holds ``BundleHeaderProto``; every other key is a variable name whose value is a
  ``BundleEntryProto`` {1 dtype, 2 shape{2 dim{1 size}}, 3 shard_id, 4 offset, 5 size, 6 crc32c}.
* ``.data-*`` holds raw little-endian row-major tensors.
"""
import os
import struct

import numpy as np

_TABLE_MAGIC = 0xDB4775248B80FB57
# tensorflow/core/framework/types.proto DataType
_DTYPES = {1: np.float32, 2: np.float64, 3: np.int32, 4: np.uint8, 5: np.int16, 6: np.int8,
           9: np.int64, 10: np.bool_}


class CheckpointError(ValueError):
    pass


def _varint(buf, pos):
    out = 0
    shift = 0
    while True:
        if pos >= len(buf):
            raise CheckpointError("truncated varint")
        b = buf[pos]
        pos += 1
        out |= (b & 0x7F) << shift
        if not b & 0x80:
            return out, pos
        shift += 7
        if shift > 63:
            raise CheckpointError("varint too long")


def _block(buf, offset, size):
    if offset + size + 5 > len(buf):
        raise CheckpointError("block handle outside file")
    if buf[offset + size] != 0:
        raise CheckpointError("compressed index block (type %d) not supported" % buf[offset + size])
    return buf[offset:offset + size]


def _block_entries(block):
    if len(block) < 4:
        raise CheckpointError("short block")
    n_restarts = struct.unpack_from("<I", block, len(block) - 4)[0]
    end = len(block) - 4 - 4 * n_restarts
    pos = 0
    key = b""
    while pos < end:
        shared, pos = _varint(block, pos)
        non_shared, pos = _varint(block, pos)
        vlen, pos = _varint(block, pos)
        key = key[:shared] + bytes(block[pos:pos + non_shared])
        pos += non_shared
        yield key, bytes(block[pos:pos + vlen])
        pos += vlen


def _proto_fields(buf):
    """Minimal protobuf wire decoder: yields (field_no, wire_type, value)."""
    pos = 0
    while pos < len(buf):
        tag, pos = _varint(buf, pos)
        field, wt = tag >> 3, tag & 7
        if wt == 0:
            v, pos = _varint(buf, pos)
        elif wt == 1:
            v = buf[pos:pos + 8]
            pos += 8
        elif wt == 2:
            ln, pos = _varint(buf, pos)
            v = buf[pos:pos + ln]
            pos += ln
        elif wt == 5:
            v = buf[pos:pos + 4]
            pos += 4
        else:
            raise CheckpointError("unsupported protobuf wire type %d" % wt)
        yield field, wt, v


def _parse_entry(value):
    dtype, shape, shard, offset, size = 0, [], 0, 0, 0
    for field, wt, v in _proto_fields(value):
        if field == 1 and wt == 0:
            dtype = v
        elif field == 2 and wt == 2:
            for f2, w2, v2 in _proto_fields(v):
                if f2 == 2 and w2 == 2:
                    dim = 0
                    for f3, w3, v3 in _proto_fields(v2):
                        if f3 == 1 and w3 == 0:
                            dim = v3
                    shape.append(dim)
        elif field == 3 and wt == 0:
            shard = v
        elif field == 4 and wt == 0:
            offset = v
        elif field == 5 and wt == 0:
            size = v
    return dtype, tuple(shape), shard, offset, size


def read_index(index_path):
    """``{variable name: (dtype_enum, shape, shard_id, offset, nbytes)}`` from a ``.index`` file."""
    with open(index_path, "rb") as fh:
        buf = fh.read()
    if len(buf) < 48:
        raise CheckpointError("index file shorter than the table footer")
    footer = buf[-48:]
    if struct.unpack("<Q", footer[40:])[0] != _TABLE_MAGIC:
        raise CheckpointError("not a TF bundle index (bad table magic)")
    pos = 0
    _, pos = _varint(footer, pos)
    _, pos = _varint(footer, pos)
    idx_off, pos = _varint(footer, pos)
    idx_size, pos = _varint(footer, pos)
    entries = {}
    for _, handle in _block_entries(_block(buf, idx_off, idx_size)):
        off, p = _varint(handle, 0)
        size, _ = _varint(handle, p)
        for key, value in _block_entries(_block(buf, off, size)):
            if key == b"":
                continue  # BundleHeaderProto
            entries[key.decode("utf-8")] = _parse_entry(value)
    return entries


def list_variables(prefix):
    """``[(name, shape)]`` like ``tf.train.list_variables``; needs only the ``.index`` file."""
    return [(k, v[1]) for k, v in sorted(read_index(prefix + ".index").items())]


def load_checkpoint(prefix, skip_optimizer=True):
    """All variables of a bundle as ``{name: np.ndarray}``."""
    entries = read_index(prefix + ".index")
    shards = {}
    out = {}
    for name, (dt, shape, shard, offset, size) in entries.items():
        if skip_optimizer and _is_optimizer_slot(name):
            continue
        if dt not in _DTYPES:
            raise CheckpointError("variable %s has unsupported dtype enum %d" % (name, dt))
        if shard not in shards:
            # the reference only ever writes single-shard bundles (Saver default)
            path = "%s.data-%05d-of-%05d" % (prefix, shard, 1)
            if not os.path.exists(path):
                raise CheckpointError("data shard %s is missing" % path)
            with open(path, "rb") as fh:
                shards[shard] = fh.read()
        raw = shards[shard][offset:offset + size]
        arr = np.frombuffer(raw, dtype=np.dtype(_DTYPES[dt]).newbyteorder("<"))
        n = int(np.prod(shape)) if shape else 1
        if arr.size != n:
            raise CheckpointError("variable %s: %d elements on disk, shape %s" % (name, arr.size, shape))
        out[name] = arr.reshape(shape).copy()
    return out


def _is_optimizer_slot(name):
    tail = name.rsplit("/", 1)[-1]
    return (tail.startswith("Adam") or "beta1_power" in name or "beta2_power" in name
            or name.startswith("apply_gradients") or "/apply_gradients/" in name
            or tail.startswith("global_step"))


def latest_checkpoint(model_dir):
    """Restates ``tf.train.latest_checkpoint`` (run_nanotensor.py:184): reads the ``checkpoint``
    text file, else picks the ``*.index`` with the largest trailing step number."""
    state = os.path.join(model_dir, "checkpoint")
    if os.path.exists(state):
        with open(state) as fh:
            for line in fh:
                if line.startswith("model_checkpoint_path:"):
                    p = line.split(":", 1)[1].strip().strip('"')
                    p = p if os.path.isabs(p) else os.path.join(model_dir, p)
                    if os.path.exists(p + ".index"):
                        return p
                    p = os.path.join(model_dir, os.path.basename(p))
                    if os.path.exists(p + ".index"):
                        return p
    best = None
    for f in os.listdir(model_dir):
        if f.endswith(".index"):
            stem = f[:-len(".index")]
            try:
                step = int(stem.rsplit("-", 1)[1])
            except (IndexError, ValueError):
                step = -1
            if best is None or step > best[0]:
                best = (step, os.path.join(model_dir, stem))
    return best[1] if best else None


# ---------------------------------------------------------------------------------------------
# name normalisation: the forward pass looks weights up by the names the CURRENT reference code
# would create (network.py), whatever generation / prefix the checkpoint was written with.
# ---------------------------------------------------------------------------------------------

def normalise_names(variables, model_name=None):
    """Return a dict keyed by canonical names.

    * strips a leading ``<model_name>/`` (run_nanotensor.py:190 wraps everything in that scope;
      the shipped test checkpoint has none),
    * ``lstm_cell/weights`` -> ``lstm_cell/kernel``, ``lstm_cell/biases`` -> ``lstm_cell/bias``
      (TF <= 1.1 names, used by both shipped checkpoints, vs TF >= 1.2),
    * drops optimizer slots and step counters.
    """
    out = {}
    for name, arr in variables.items():
        if _is_optimizer_slot(name):
            continue
        n = name
        if model_name and n.startswith(model_name + "/"):
            n = n[len(model_name) + 1:]
        if n.endswith("lstm_cell/weights"):
            n = n[:-len("weights")] + "kernel"
        elif n.endswith("lstm_cell/biases"):
            n = n[:-len("biases")] + "bias"
        out[n] = arr
    return out


def _guess_prefix(names, network):
    """If every variable shares one leading scope that is not a layer name, treat it as the
    ``model_name`` scope and strip it."""
    layer_names = {l.get("name") for l in network}
    heads = {n.split("/", 1)[0] for n in names if "/" in n}
    top = {n for n in names if "/" not in n}
    if len(heads) == 1 and not top:
        h = next(iter(heads))
        if h not in layer_names:
            return h
    return None


def weights_for_network(variables, network, model_name=None):
    """Map checkpoint variables onto the layer list of a config (network.py:128-187).

    Returns ``{"layers": [per-layer dict of arrays], "out_w": [F,C], "out_b": [C]}`` with the TF
    layouts kept as they are on disk (kernel ``[D+H, 4H]`` gate blocks i|j|f|o; conv filters
    ``[1,k,Cin,Cout]``).  Missing variables raise ``KeyError`` naming the variable.
    """
    if model_name is None:
        model_name = _guess_prefix([n for n in variables if not _is_optimizer_slot(n)], network)
    v = normalise_names(variables, model_name)

    def get(name):
        if name not in v:
            raise KeyError("checkpoint has no variable %r (have: %s)" % (name, sorted(v)[:12]))
        return np.ascontiguousarray(v[name], dtype=np.float32)

    layers = []
    for layer in network:
        t, nm = layer["type"], layer["name"]
        if t == "blstm":
            base = nm + "/bidirectional_rnn/"
            layers.append({
                "kernel_fw": get(base + "fw/lstm_cell/kernel"), "bias_fw": get(base + "fw/lstm_cell/bias"),
                "kernel_bw": get(base + "bw/lstm_cell/kernel"), "bias_bw": get(base + "bw/lstm_cell/bias")})
        elif t == "residual_layer":
            d = {"conv1": get(nm + "/branch1/conv1/weights")}
            if layer.get("batchnorm"):
                d["conv1_bn_scale"] = get(nm + "/branch1/conv1_bn/conv1_bn_scale")
                d["conv1_bn_offset"] = get(nm + "/branch1/conv1_bn/conv1_bn_offset")
            for c in ("conv2a", "conv2b", "conv2c"):
                d[c] = get("%s/branch2/%s/weights" % (nm, c))
                d[c + "_bn_scale"] = get("%s/branch2/%s_bn/%s_bn_scale" % (nm, c, c))
                d[c + "_bn_offset"] = get("%s/branch2/%s_bn/%s_bn_offset" % (nm, c, c))
            layers.append(d)
        elif t == "chiron_fnn":
            layers.append({"weights": get(nm + "/weights"), "bias": get(nm + "/bias")})
        else:
            raise KeyError("layer type %r is not on the supported forward path" % t)
    # output layer variables are created under a name_scope only -> un-scoped names
    # (network.py:79-80, 298-301)
    return {"layers": layers, "out_w": get("weights"), "out_b": get("bias")}
