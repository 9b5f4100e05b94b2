"""Executor objects with the reference's backend-plugin interface.

The reference plugs an accelerated sub-graph in as a ``ModelExectuor`` (sic — the reference's spelling, kept so the
type can be used interchangeably; ref: sam2/ytools/executor/model_executor.py:110-169) stored in
``backend_contexts`` and called as ``Inference([tensors...], output_type="torch")``.  ``B200Executor`` is that object
for the sm_100a engine: same methods, same I/O tensor names as the reference's exported ONNX graphs
(ref: export_image_onnx.py:49-77,128-148; export_memory_attention_onnx.py:52-90; export_other_onnx.py:151-195),
same error behaviour (wrong input count -> AssertionError, bad output_type -> Exception).
"""
import threading
from abc import ABC, abstractmethod
from typing import Dict, List, Optional, Tuple

import numpy as np
import torch


class TensorDataType:
    FLOAT32 = "float32"
    BFLOAT16 = "bfloat16"


class TensorDesc:
    def __init__(self, name: str, shape: Tuple[int], dtype=TensorDataType.FLOAT32, dynamic_desc: dict = None):
        self.name, self.shape, self.dtype = name, shape, dtype
        self.dynamic_desc = dynamic_desc if dynamic_desc is not None else dict()

    def __str__(self):
        s = f'name="{self.name}", shape={self.shape}, dtype={self.dtype}'
        return s + (f", dynamic_desc={self.dynamic_desc}" if self.dynamic_desc else "")


class ModelExectuor(ABC):
    def warmup(self, inputs, times=3):
        for _ in range(times):
            self.Inference(inputs)

    def SetInputShapes(self, shapes):
        pass

    @abstractmethod
    def Inference(self, inputs: list, output_type="numpy") -> list:
        return []

    @abstractmethod
    def GetModelInputDesc(self) -> List[TensorDesc]:
        return []

    @abstractmethod
    def GetModelOutputDesc(self) -> List[TensorDesc]:
        return []

    def PrintIODesc(self):
        print("input desc:")
        for d in self.GetModelInputDesc():
            print(d)
        print("output desc:")
        for d in self.GetModelOutputDesc():
            print(d)

    @abstractmethod
    def Release(self):
        pass


_N = {0: "N"}
_IO = {
    # kind: (inputs, outputs)
    "set_image_e2e": ([("image", (-1, 3, 1024, 1024), _N)],
                      [("feature0", (-1, 32, 256, 256), _N), ("feature1", (-1, 64, 128, 128), _N), ("feature2", (-1, 256, 64, 64), _N)]),
    "inference_image": ([("image", (-1, 3, 1024, 1024), _N)],
                        [("vision_features", (-1, 256, 64, 64), _N), ("vision_pos_enc0", (-1, 256, 256, 256), _N),
                         ("vision_pos_enc1", (-1, 256, 128, 128), _N), ("vision_pos_enc2", (-1, 256, 64, 64), _N),
                         ("backbone_fpn0", (-1, 32, 256, 256), _N), ("backbone_fpn1", (-1, 64, 128, 128), _N),
                         ("backbone_fpn2", (-1, 256, 64, 64), _N)]),
    "memory_attention_exclude": ([("curr", (4096, -1, 256), {1: "N"}), ("memory", (-1, 4096, -1, 64), {0: "L", 2: "N"}),
                                  ("curr_pos", (4096, -1, 256), {1: "N"}), ("memory_pos", (-1, 4096, -1, 64), {0: "L", 2: "N"}),
                                  ("memory_exclude", (-1, -1, 64), {0: "P", 1: "N"}), ("memory_pos_exclude", (-1, -1, 64), {0: "P", 1: "N"})],
                                 [("memory_output", (4096, -1, 256), {1: "N"})]),
    "memory_encoder": ([("pixel_features", (-1, 256, 64, 64), _N), ("mask_for_memory", (-1, 1, 1024, 1024), _N)],
                       [("mask_memory_features", (-1, 64, 64, 64), _N), ("mask_memory_pos_enc", (-1, 64, 64, 64), _N)]),
}
# the reference's `none` engine drops the two unused pointer inputs (memory_attention.py:376-378)
_IO["memory_attention_none"] = (_IO["memory_attention_exclude"][0][:4], _IO["memory_attention_exclude"][1])


class B200Executor(ModelExectuor):
    """One accelerated sub-graph on a shared ``Engine``.  Inputs may be torch tensors (any device) or numpy arrays;
    they are moved to the engine's device if needed (the reference stages every input too, ytools/tensorrt/memory.py:64-79).
    Outputs are new fp32 CUDA tensors owned by the caller (or numpy arrays for output_type="numpy")."""

    def __init__(self, engine, kind: str):
        if kind not in _IO:
            raise ValueError(f"unknown sub-graph kind {kind!r}; expected one of {sorted(_IO)}")
        self.engine, self.kind = engine, kind
        self._lock = threading.Lock()
        self._released = False

    # -- descriptors ---------------------------------------------------------------------------------------
    def GetModelInputDesc(self) -> List[TensorDesc]:
        return [TensorDesc(n, s, TensorDataType.FLOAT32, dict(d)) for n, s, d in _IO[self.kind][0]]

    def GetModelOutputDesc(self) -> List[TensorDesc]:
        return [TensorDesc(n, s, TensorDataType.FLOAT32, dict(d)) for n, s, d in _IO[self.kind][1]]

    # -- run -------------------------------------------------------------------------------------------------
    def _to_dev(self, x):
        if isinstance(x, np.ndarray):
            x = torch.from_numpy(x)
        return x.to(self.engine.device, torch.float32)

    def Inference(self, inputs: list, output_type="numpy") -> list:
        if self._released:
            raise RuntimeError("B200Executor used after Release()")
        n_in = len(_IO[self.kind][0])
        assert len(inputs) == n_in, f"{self.kind}: expected {n_in} inputs, got {len(inputs)}"
        if output_type not in ("numpy", "torch"):
            raise Exception(f"unsupported output_type={output_type}")
        with self._lock:
            x = [self._to_dev(t) for t in inputs]
            e = self.engine
            if self.kind == "set_image_e2e":
                outs = list(e.image_encode(x[0], normalised=False, add_no_mem_embed=True))
            elif self.kind == "inference_image":
                f0, f1, f2 = e.image_encode(x[0], normalised=True, add_no_mem_embed=False)
                n = x[0].shape[0]
                outs = [f2, e.position_encoding(256, 256, n), e.position_encoding(256, 128, n), e.position_encoding(256, 64, n), f0, f1, f2]
            elif self.kind == "memory_attention_none":
                outs = [e.memory_attention(x[0], x[1], x[2], x[3])]
            elif self.kind == "memory_attention_exclude":
                P = x[4].shape[0]
                outs = [e.memory_attention(x[0], x[1], x[2], x[3], x[4] if P else None, x[5] if P else None)]
            else:
                outs = list(e.memory_encode(x[0], x[1]))
        if output_type == "numpy":
            return [o.cpu().numpy() for o in outs]
        return outs

    def Release(self):
        """Idempotent; the engine itself is shared and released by whoever created it."""
        self._released = True
        self.engine = None
