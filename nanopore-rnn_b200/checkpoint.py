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


# ---------------------------------------------------------------------------------------------
# writer: emits the same bundle format (single shard, uncompressed blocks, masked CRC32C), so
# synthetic / converted weights can be handed to the unmodified config keys
# (trained_model_path / trained_model + use_checkpoint).
# ---------------------------------------------------------------------------------------------
_CRC_TABLE = None


def _crc32c(data, crc=0):
    global _CRC_TABLE
    if _CRC_TABLE is None:
        tab = []
        for i in range(256):
            c = i
            for _ in range(8):
                c = (c >> 1) ^ 0x82F63B78 if c & 1 else c >> 1
            tab.append(c)
        _CRC_TABLE = np.array(tab, dtype=np.uint32)
    crc ^= 0xFFFFFFFF
    tab = _CRC_TABLE
    for b in bytes(data):
        crc = int(tab[(crc ^ b) & 0xFF]) ^ (crc >> 8)
    return crc ^ 0xFFFFFFFF


def _mask_crc(crc):
    return (((crc >> 15) | (crc << 17)) + 0xA282EAD8) & 0xFFFFFFFF


def _enc_varint(v):
    out = bytearray()
    while True:
        b = v & 0x7F
        v >>= 7
        if v:
            out.append(b | 0x80)
        else:
            out.append(b)
            return bytes(out)


def _enc_field(field, wt, payload):
    return _enc_varint((field << 3) | wt) + payload


def _enc_entry(dtype_enum, shape, offset, size, crc):
    shape_msg = b"".join(_enc_field(2, 2, _enc_varint(len(d)) + d)
                         for d in (_enc_field(1, 0, _enc_varint(int(s))) for s in shape))
    msg = _enc_field(1, 0, _enc_varint(dtype_enum))
    msg += _enc_field(2, 2, _enc_varint(len(shape_msg)) + shape_msg)
    if offset:
        msg += _enc_field(4, 0, _enc_varint(offset))
    msg += _enc_field(5, 0, _enc_varint(size))
    msg += _enc_field(6, 5, struct.pack("<I", _mask_crc(crc)))
    return msg


def _build_block(items):
    """items: [(key bytes, value bytes)] sorted; no prefix sharing, one restart per entry."""
    body = bytearray()
    restarts = []
    for k, v in items:
        restarts.append(len(body))
        body += _enc_varint(0) + _enc_varint(len(k)) + _enc_varint(len(v)) + k + v
    if not restarts:
        restarts = [0]
    for r in restarts:
        body += struct.pack("<I", r)
    body += struct.pack("<I", len(restarts))
    return bytes(body)


def save_checkpoint(prefix, variables):
    """Write ``{name: array}`` as ``<prefix>.index`` + ``<prefix>.data-00000-of-00001``."""
    enum_of = {np.dtype(v): k for k, v in _DTYPES.items()}
    data = bytearray()
    items = []
    for name in sorted(variables, key=lambda s: s.encode("utf-8")):
        arr = np.ascontiguousarray(variables[name])
        if arr.dtype not in enum_of:
            raise CheckpointError("variable %s: dtype %s not supported" % (name, arr.dtype))
        raw = arr.astype(arr.dtype.newbyteorder("<"), copy=False).tobytes()
        items.append((name.encode("utf-8"), _enc_entry(enum_of[arr.dtype], arr.shape, len(data), len(raw),
                                                       _crc32c(raw))))
        data += raw
    # BundleHeaderProto {1: num_shards = 1, 2: endianness = LITTLE(0), 3: version{1: producer = 1}}
    header = _enc_field(1, 0, _enc_varint(1)) + _enc_field(3, 2, _enc_varint(2) + _enc_field(1, 0, _enc_varint(1)))
    items = [(b"", header)] + items
    out = bytearray()

    def emit(block):
        off = len(out)
        out.extend(block)
        trailer_crc = _mask_crc(_crc32c(b"\x00", _crc32c(block)))
        out.extend(b"\x00" + struct.pack("<I", trailer_crc))
        return off, len(block)

    d_off, d_size = emit(_build_block(items))
    m_off, m_size = emit(_build_block([]))
    handle = _enc_varint(d_off) + _enc_varint(d_size)
    i_off, i_size = emit(_build_block([(items[-1][0] + b"\x00", handle)]))
    footer = _enc_varint(m_off) + _enc_varint(m_size) + _enc_varint(i_off) + _enc_varint(i_size)
    footer += b"\x00" * (40 - len(footer)) + struct.pack("<Q", _TABLE_MAGIC)
    out.extend(footer)
    with open(prefix + ".index", "wb") as fh:
        fh.write(bytes(out))
    with open(prefix + ".data-00000-of-00001", "wb") as fh:
        fh.write(bytes(data))
    with open(os.path.join(os.path.dirname(prefix) or ".", "checkpoint"), "w") as fh:
        fh.write('model_checkpoint_path: "%s"\n' % os.path.basename(prefix))
    return prefix


def variables_from_weights(weights, network, model_name=None, old_names=False):
    """Inverse of ``weights_for_network``: the TF variable names the reference's code creates
    (network.py:165-172, 298-301, 315-347 and rnn.LSTMCell's kernel/bias)."""
    pre = model_name + "/" if model_name else ""
    kern, bias = ("weights", "biases") if old_names else ("kernel", "bias")
    out = {}
    for layer, w in zip(network, weights["layers"]):
        nm = pre + layer["name"]
        if layer["type"] == "blstm":
            for d in ("fw", "bw"):
                out["%s/bidirectional_rnn/%s/lstm_cell/%s" % (nm, d, kern)] = w["kernel_" + d]
                out["%s/bidirectional_rnn/%s/lstm_cell/%s" % (nm, d, bias)] = w["bias_" + d]
        elif layer["type"] == "residual_layer":
            out[nm + "/branch1/conv1/weights"] = w["conv1"]
            if layer.get("batchnorm"):
                out[nm + "/branch1/conv1_bn/conv1_bn_scale"] = w["conv1_bn_scale"]
                out[nm + "/branch1/conv1_bn/conv1_bn_offset"] = w["conv1_bn_offset"]
            for c in ("conv2a", "conv2b", "conv2c"):
                out["%s/branch2/%s/weights" % (nm, c)] = w[c]
                out["%s/branch2/%s_bn/%s_bn_scale" % (nm, c, c)] = w[c + "_bn_scale"]
                out["%s/branch2/%s_bn/%s_bn_offset" % (nm, c, c)] = w[c + "_bn_offset"]
        elif layer["type"] == "chiron_fnn":
            out[nm + "/weights"] = w["weights"]
            out[nm + "/bias"] = w["bias"]
    out[pre + "weights"] = weights["out_w"]
    out[pre + "bias"] = weights["out_b"]
    out[pre + "global_step"] = np.float32(0.0).reshape(())
    return out
