"""Python mirror of the host-side model loading (vox_load,
/root/reference/voxtral.c:131-184): mmap consolidated.safetensors, hand bf16
matrices to the CUDA backend as pointers into the mapping, convert the small
tensors to f32, precompute ada_scale (voxtral.c:31-80).  Used by bench.py and
the tests; the C host in host/ does the same for the CLI.

A Q8 container (dtype "Q8" = [rows x f32 scale][rows x cols int8], /root/reference/quantize.py:111-129,
loader voxtral_safetensors.c:393-410,453-468) is detected from the file: the matrices go to the backend as
(int8 rows, scales) pointer pairs (vox_cuda_weights_q8_t), everything else is F32 in such a file."""
import ctypes as C
import json
import mmap
import os
import struct

import numpy as np

import voxcuda as vc

ENC = "mm_streams_embeddings.embedding_module.whisper_encoder"


def bf16_to_f32(u16):
    return (u16.astype(np.uint32) << 16).view(np.float32)


def gelu_tanh(x):
    x = x.astype(np.float32)
    inner = np.float32(0.7978845608028654) * (x + np.float32(0.044715) * x * x * x)
    return np.float32(0.5) * x * (np.float32(1.0) + np.tanh(inner))


class HostModel:
    def __init__(self, model_dir, delay_tokens=6):
        path = os.path.join(model_dir, "consolidated.safetensors")
        self.f = open(path, "rb")
        self.mm = mmap.mmap(self.f.fileno(), 0, access=mmap.ACCESS_READ)
        (hlen,) = struct.unpack("<Q", self.mm[:8])
        self.header = json.loads(self.mm[8:8 + hlen])
        self.header.pop("__metadata__", None)
        self.base = 8 + hlen
        self.raw = np.frombuffer(self.mm, dtype=np.uint8)
        self.keep = []
        self.n_enc = sum(1 for k in self.header if k.endswith("attention.wq.weight") and k.startswith(ENC))
        self.n_dec = sum(1 for k in self.header if k.endswith("attention.wq.weight") and k.startswith("layers."))
        self.q8 = self.header["layers.0.attention.wq.weight"]["dtype"] == "Q8"      # ctx->use_q8, voxtral.c:175-176
        self.weights_q8 = vc.WeightsQ8() if self.q8 else None
        self.weights = self._build(delay_tokens)

    def create_model(self, L):
        """vox_cuda_model_create / _create_q8, whichever the file calls for"""
        if self.q8:
            return L.vox_cuda_model_create_q8(C.byref(self.weights), C.byref(self.weights_q8))
        return L.vox_cuda_model_create(C.byref(self.weights))

    def u16(self, name):
        t = self.header[name]
        assert t["dtype"] == "BF16", name
        a, b = t["data_offsets"]
        return self.raw[self.base + a:self.base + b].view(np.uint16)

    def f32(self, name):
        t = self.header[name]
        if t["dtype"] == "F32":
            a, b = t["data_offsets"]
            arr = np.ascontiguousarray(self.raw[self.base + a:self.base + b].view(np.float32))
        else:
            arr = np.ascontiguousarray(bf16_to_f32(self.u16(name)))
        self.keep.append(arr)
        return arr

    def q8_parts(self, name):
        """(int8 [rows, cols], f32 scales [rows]) views of a Q8 tensor"""
        t = self.header[name]
        assert t["dtype"] == "Q8", name
        rows, cols = t["shape"]
        a, b = t["data_offsets"]
        assert b - a == rows * 4 + rows * cols
        sc = self.raw[self.base + a:self.base + a + rows * 4].view(np.float32)
        q = self.raw[self.base + a + rows * 4:self.base + b].view(np.int8).reshape(rows, cols)
        return q, sc

    def dense_f32(self, name):
        """a small matrix as fp32, whatever the container (safetensors_get_f32, voxtral_safetensors.c:362-410)"""
        t = self.header[name]
        if t["dtype"] == "Q8":
            q, sc = self.q8_parts(name)
            return q.astype(np.float32) * sc[:, None]
        if t["dtype"] == "F32":
            return self.f32(name).reshape(t["shape"])
        return bf16_to_f32(self.u16(name)).reshape(t["shape"])

    def _mat(self, holder, field, holder_q8, name):
        """bind one 2-D matrix: bf16 pointer, or the (int8, scale) pair in the Q8 table"""
        if self.q8:
            q, sc = self.q8_parts(name)
            setattr(holder_q8, field, vc.Q8(C.cast(q.ctypes.data, vc.i8p), C.cast(sc.ctypes.data, vc.f32p)))
        else:
            setattr(holder, field, self._p16(name))

    def _p16(self, name):
        a = self.u16(name)
        return C.cast(a.ctypes.data, vc.u16p)

    def _p32(self, name):
        return self.f32(name).ctypes.data_as(vc.f32p)

    def ada_scale(self, delay_tokens):
        half = 1536
        i = np.arange(half, dtype=np.float32)
        inv_freq = np.exp(-np.float32(np.log(np.float32(10000.0))) * i / np.float32(half)).astype(np.float32)
        emb = np.float32(delay_tokens) * inv_freq
        t_cond = np.concatenate([np.cos(emb), np.sin(emb)]).astype(np.float32)
        out = np.empty((self.n_dec, 3072), np.float32)
        for l in range(self.n_dec):
            down = self.dense_f32(f"layers.{l}.ada_rms_norm_t_cond.0.weight").reshape(32, 3072)
            up = self.dense_f32(f"layers.{l}.ada_rms_norm_t_cond.2.weight").reshape(3072, 32)
            out[l] = up @ gelu_tanh(down @ t_cond)
        return out

    def _build(self, delay_tokens):
        w = vc.Weights()
        q = self.weights_q8
        w.n_enc_layers, w.n_dec_layers = self.n_enc, self.n_dec
        w.conv0_weight = self._p32(f"{ENC}.conv_layers.0.conv.weight")
        w.conv0_bias = self._p32(f"{ENC}.conv_layers.0.conv.bias")
        w.conv1_weight = self._p32(f"{ENC}.conv_layers.1.conv.weight")
        w.conv1_bias = self._p32(f"{ENC}.conv_layers.1.conv.bias")
        for i in range(self.n_enc):
            p, d = f"{ENC}.transformer.layers.{i}.", w.enc[i]
            dq = q.enc[i] if q is not None else None
            for n in ("wq", "wk", "wv", "wo"):
                self._mat(d, n, dq, p + f"attention.{n}.weight")
            for n in ("w1", "w2", "w3"):
                self._mat(d, n, dq, p + f"feed_forward.{n}.weight")
            d.wq_bias = self._p32(p + "attention.wq.bias")
            d.wv_bias = self._p32(p + "attention.wv.bias")
            d.wo_bias = self._p32(p + "attention.wo.bias")
            d.w2_bias = self._p32(p + "feed_forward.w2.bias")
            d.attention_norm = self._p32(p + "attention_norm.weight")
            d.ffn_norm = self._p32(p + "ffn_norm.weight")
        w.enc_norm = self._p32(f"{ENC}.transformer.norm.weight")
        self._mat(w, "adapter_w0", q, "mm_streams_embeddings.embedding_module.audio_language_projection.0.weight")
        self._mat(w, "adapter_w1", q, "mm_streams_embeddings.embedding_module.audio_language_projection.2.weight")
        self._mat(w, "tok_embeddings", q, "mm_streams_embeddings.embedding_module.tok_embeddings.weight")
        for i in range(self.n_dec):
            p, d = f"layers.{i}.", w.dec[i]
            dq = q.dec[i] if q is not None else None
            for n in ("wq", "wk", "wv", "wo"):
                self._mat(d, n, dq, p + f"attention.{n}.weight")
            for n in ("w1", "w2", "w3"):
                self._mat(d, n, dq, p + f"feed_forward.{n}.weight")
            d.attention_norm = self._p32(p + "attention_norm.weight")
            d.ffn_norm = self._p32(p + "ffn_norm.weight")
        w.dec_norm = self._p32("norm.weight")
        self.ada = np.ascontiguousarray(self.ada_scale(delay_tokens))
        w.ada_scale = self.ada.ctypes.data_as(vc.f32p)
        return w
