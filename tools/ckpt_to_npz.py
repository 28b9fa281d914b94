"""Offline converter: MindSpore `.ckpt` (protobuf) -> `.npz` with the same parameter names.

    python tools/ckpt_to_npz.py models/wukong-huahua-ms.ckpt            # -> models/wukong-huahua-ms.npz
    python tools/ckpt_to_npz.py lora.ckpt --out lora.npz --list

`minddiffusion_b200.txt2img` and the model classes load weights from `{MindSpore parameter name: array}`; this tool
produces that mapping without MindSpore or the protobuf runtime.  The wire format is restated from MindSpore's
`checkpoint.proto` (mindspore/ccsrc/utils/checkpoint.proto, MindSpore 1.x / 2.x):

    message Checkpoint  { repeated Value value = 1; }
    message Value       { required string tag = 1; optional TensorProto tensor = 2; }
    message TensorProto { repeated int64 dims = 1; required string tensor_type = 2; required bytes tensor_content = 3; }

`save_checkpoint` splits a tensor above its slice size into consecutive `Value`s with the same tag; they are
concatenated here as `load_checkpoint` does.  NOT verified against a MindSpore-written file in the build container
(none is available there and MindSpore cannot be installed); `tests/test_cpu.py` checks the parser against a writer
of the same schema (round trip), which pins the varint / length-delimited handling but not MindSpore itself.
"""
import argparse
import os
import sys

import numpy as np

DTYPES = {"Float32": np.float32, "Float16": np.float16, "Float64": np.float64, "Int8": np.int8, "Int16": np.int16,
          "Int32": np.int32, "Int64": np.int64, "UInt8": np.uint8, "UInt16": np.uint16, "UInt32": np.uint32,
          "UInt64": np.uint64, "Bool": np.bool_}


def _varint(buf, pos):
    shift, val = 0, 0
    while True:
        b = buf[pos]
        pos += 1
        val |= (b & 0x7F) << shift
        if not b & 0x80:
            return val, pos
        shift += 7
        if shift > 70:
            raise ValueError("malformed varint at byte %d" % pos)


def _fields(buf):
    """Yield (field number, wire type, value) of one message; length-delimited values are memoryviews."""
    pos, end = 0, len(buf)
    while pos < end:
        key, pos = _varint(buf, pos)
        num, wt = key >> 3, key & 7
        if wt == 0:
            val, pos = _varint(buf, pos)
        elif wt == 2:
            n, pos = _varint(buf, pos)
            val = buf[pos:pos + n]
            pos += n
        elif wt == 1:
            val = bytes(buf[pos:pos + 8])
            pos += 8
        elif wt == 5:
            val = bytes(buf[pos:pos + 4])
            pos += 4
        else:
            raise ValueError("unsupported wire type %d at byte %d" % (wt, pos))
        yield num, wt, val


def _signed(v):
    return v - (1 << 64) if v >= (1 << 63) else v


def parse_checkpoint(data):
    """bytes -> {tag: ndarray} (insertion order = file order)."""
    buf = memoryview(data)
    raw = {}
    for num, wt, val in _fields(buf):
        if num != 1 or wt != 2:
            continue
        tag, tensor = None, None
        for n2, w2, v2 in _fields(val):
            if n2 == 1 and w2 == 2:
                tag = bytes(v2).decode("utf-8")
            elif n2 == 2 and w2 == 2:
                tensor = v2
        if tag is None or tensor is None:
            continue  # map tensors and other value kinds carry no dense weights
        dims, ttype, content = [], None, b""
        for n3, w3, v3 in _fields(tensor):
            if n3 == 1 and w3 == 0:
                dims.append(_signed(v3))
            elif n3 == 1 and w3 == 2:  # packed encoding of the repeated int64
                p = 0
                while p < len(v3):
                    d, p = _varint(v3, p)
                    dims.append(_signed(d))
            elif n3 == 2 and w3 == 2:
                ttype = bytes(v3).decode("utf-8")
            elif n3 == 3 and w3 == 2:
                content = bytes(v3)
        if tag in raw:  # a sliced tensor: consecutive values with one tag
            raw[tag][2].append(content)
        else:
            raw[tag] = (dims, ttype, [content])
    out = {}
    for tag, (dims, ttype, chunks) in raw.items():
        if ttype == "BFloat16":  # numpy has no bfloat16: widen to float32
            u16 = np.frombuffer(b"".join(chunks), dtype=np.uint16)
            arr = (u16.astype(np.uint32) << 16).view(np.float32)
        elif ttype in DTYPES:
            arr = np.frombuffer(b"".join(chunks), dtype=DTYPES[ttype])
        else:
            raise ValueError("%s: unsupported tensor_type %r" % (tag, ttype))
        shape = tuple(int(d) for d in dims)
        if int(np.prod(shape)) != arr.size:
            raise ValueError("%s: %d elements do not fill shape %s" % (tag, arr.size, shape))
        out[tag] = arr.reshape(shape)
    return out


# ---- writer of the same schema (used by the round-trip test; also handy to hand weights back to MindSpore) ----
def _enc_varint(v):
    v &= (1 << 64) - 1
    out = bytearray()
    while True:
        b = v & 0x7F
        v >>= 7
        out.append(b | (0x80 if v else 0))
        if not v:
            return bytes(out)


def _ld(num, payload):
    return _enc_varint((num << 3) | 2) + _enc_varint(len(payload)) + payload


def write_checkpoint(params, slice_bytes=None):
    names = {np.dtype(v): k for k, v in DTYPES.items()}
    blob = bytearray()
    for tag, arr in params.items():
        arr = np.asarray(arr)
        content = arr.tobytes()  # C order; 0-d arrays keep their empty dims
        step = slice_bytes or max(len(content), 1)
        for off in range(0, max(len(content), 1), step):
            tensor = b"".join(_enc_varint((1 << 3) | 0) + _enc_varint(d) for d in arr.shape)
            tensor += _ld(2, names[arr.dtype].encode()) + _ld(3, content[off:off + step])
            blob += _ld(1, _ld(1, tag.encode("utf-8")) + _ld(2, tensor))
    return bytes(blob)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("ckpt")
    ap.add_argument("--out", default=None)
    ap.add_argument("--list", action="store_true", help="print name, dtype, shape of every entry")
    ap.add_argument("--keep-dtype", action="store_true", help="keep fp16 weights as fp16 (default: widen to float32)")
    a = ap.parse_args()
    params = parse_checkpoint(open(a.ckpt, "rb").read())
    if not a.keep_dtype:
        params = {k: (v.astype(np.float32) if v.dtype == np.float16 else v) for k, v in params.items()}
    if a.list:
        for k, v in params.items():
            print("%-90s %-8s %s" % (k, v.dtype, tuple(v.shape)))
    out = a.out or os.path.splitext(a.ckpt)[0] + ".npz"
    np.savez(out, **params)
    print("wrote %s (%d tensors, %.1f M values)" % (out, len(params), sum(v.size for v in params.values()) / 1e6),
          file=sys.stderr)


if __name__ == "__main__":
    main()
