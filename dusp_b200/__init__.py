"""dusp_b200 -- B200-native batch renderer for Dusp circuits.

Host-side mirror of the reference's public surface (``src/index.js:1-20``):
``unDusp``, ``renderChannelData``, ``RenderStream``, ``quick``, ``Unit``, ``Patch``, ``Circuit``,
``components``, ``patches``.  Graph construction is host logic; every sample is
computed by the CUDA library behind ``include/dusp_b200.h``.
"""
from . import config
from .graph import Unit, Patch, Circuit, Inlet, Outlet, DuspError, Event
from . import components, patches, quick
from .render import renderChannelData, render_batch, Batch
from .construct import unDusp
from .stream import RenderStream

__all__ = ["config", "Unit", "Patch", "Circuit", "Inlet", "Outlet", "DuspError", "Event", "components",
           "patches", "quick", "renderChannelData", "render_batch", "Batch", "unDusp", "RenderStream"]
