"""ReSTIR DI trace pass: host mirror of render::rtdi::Trace (src/render/rtdi/Trace.hpp:19-56, Trace.cpp:
176-362) and of the pass that feeds it, rtdi::InitialReservoirs (src/render/rtdi/InitialReservoirs.cpp). Same names, same
argument meaning, the same history / frame-index behaviour; Vulkan image handles become numpy arrays or device images held
by the context. The raster G-buffer pass is out of scope: `GBuffer.from_primary_hits` asks the tracer for the texels a raster
pass would have written (prt_rtdi_gbuffer_from_primary), or the caller uploads its own G-buffer."""
from __future__ import annotations

import ctypes as C
import dataclasses
from typing import Optional

import numpy as np

from . import rt_reference as rt
from . import scene_data as sd


@dataclasses.dataclass
class GBuffer:
    """GBufferRendererOutput as the trace pass reads it (formats GBufferRenderer.cpp:31-32)."""
    albedoRoughness: np.ndarray  # uint32 [h, w], R8G8B8A8_UNORM
    normalMetalness: np.ndarray  # uint32 [h, w], A2B10G10R10_UNORM_PACK32
    depth: np.ndarray            # float32 [h, w], non-linear reverse-Z, 0 = nothing drawn

    def arrays(self):
        return self.albedoRoughness, self.normalMetalness, self.depth

    @staticmethod
    def from_primary_hits(world: rt.World, width: int, height: int, frame_index: int = 1) -> "GBuffer":
        rt._check(world.lib.prt_rtdi_gbuffer_from_primary(world.handle, width, height, frame_index))
        ar, nm = np.zeros((height, width), np.uint32), np.zeros((height, width), np.uint32)
        z = np.zeros((height, width), np.float32)
        rt._check(world.lib.prt_rtdi_read_gbuffer(world.handle, rt._p(ar), rt._p(nm), rt._p(z)))
        return GBuffer(ar, nm, z)


class InitialReservoirs:
    """rtdi::InitialReservoirs: restir_di/initial_reservoirs.comp over the G-buffer the context holds."""

    def __init__(self):
        self.m_frameIndex = 0

    def record(self, world: rt.World, width: int, height: int) -> np.ndarray:
        self.m_frameIndex = (self.m_frameIndex + 1) % 4096  # InitialReservoirs.cpp: the same 4096 period as the trace
        rt._check(world.lib.prt_rtdi_initial_reservoirs(world.handle, self.m_frameIndex))
        out = np.zeros((height, width, 2), np.float32)
        rt._check(world.lib.prt_rtdi_read_reservoirs(world.handle, rt._p(out)))
        return out


@dataclasses.dataclass
class TraceInput:  # Trace::Input (Trace.hpp:40-44)
    gbuffer: Optional[GBuffer]        # None: keep the G-buffer the context already holds (from_primary_hits)
    reservoirs: Optional[np.ndarray]  # float32 [h, w, 2]; None: keep the reservoirs the context already holds


class TraceOutput:  # Trace::Output (Trace.hpp:45-48)
    def __init__(self, world: rt.World, width: int, height: int):
        self._world, self.width, self.height = world, width, height

    def illumination(self) -> np.ndarray:
        out = np.zeros((self.height, self.width, 4), np.float16)
        rt._check(self._world.lib.prt_rtdi_read_illumination_f16(self._world.handle, rt._p(out), out.nbytes))
        return out

    def illumination_f32(self) -> np.ndarray:
        out = np.zeros((self.height, self.width, 4), np.float32)
        rt._check(self._world.lib.prt_rtdi_read_output_f32(self._world.handle, rt._p(out), out.nbytes))
        return out


class PrtRtdiTracePC(C.Structure):
    _fields_ = [("drawType", C.c_uint32), ("frameIndex", C.c_uint32), ("flags", C.c_uint32)]


class Trace:
    """render::rtdi::Trace."""

    def __init__(self):
        self.m_initialized = False
        self.m_accumulationDirty = True
        self.m_accumulate = True
        self.m_frameIndex = 0

    def init(self, world: rt.World):
        if self.m_initialized:
            raise RuntimeError("Trace::init: already initialized")
        self.m_initialized = True

    def recompileShaders(self) -> bool:
        return True  # nothing to compile at run time

    def record(self, world: rt.World, width: int, height: int, input: TraceInput, resetAccumulation: bool = False, drawType: int = 0,
               cameraChanged: bool = False) -> TraceOutput:
        if not self.m_initialized:
            raise RuntimeError("Trace::record: not initialized")
        self.m_frameIndex = (self.m_frameIndex + 1) % 4096  # Trace.cpp:184
        if input.gbuffer is not None:
            ar, nm, z = (np.ascontiguousarray(a) for a in input.gbuffer.arrays())
            if ar.shape != (height, width):
                raise RuntimeError("G-buffer extent differs from the render extent")
            rt._check(world.lib.prt_rtdi_set_gbuffer(world.handle, width, height, rt._p(ar), rt._p(nm), rt._p(z)))
        if input.reservoirs is not None:
            rt._check(world.lib.prt_rtdi_set_reservoirs(world.handle, rt._p(np.ascontiguousarray(input.reservoirs, np.float32))))
        skip = cameraChanged or resetAccumulation or self.m_accumulationDirty  # Trace.cpp:255-263
        pc = PrtRtdiTracePC(drawType, self.m_frameIndex, (sd.FLAG_SKIP_HISTORY if skip else 0) | (sd.FLAG_ACCUMULATE if self.m_accumulate else 0))
        rt._check(world.lib.prt_rtdi_trace(world.handle, C.byref(pc)))
        self.m_accumulationDirty = False
        return TraceOutput(world, width, height)

    def releasePreserved(self, world: rt.World):
        rt._check(world.lib.prt_rtdi_release_preserved(world.handle))
