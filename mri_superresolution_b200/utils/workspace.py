"""Caller-owned workspaces of the C ABI (include/mrisr.h: zero-filled once per shape, then reused), cached per shape.

A recorded CUDA graph (utils/graph.py) holds raw pointers into the workspace it was captured with, so that tensor must
outlive the graph no matter which other shapes the module sees in between (a ragged last batch of an epoch run eagerly, a
validation pass under ``no_grad`` that asks for the forward-only workspace, ...).  ``WorkspaceCache`` therefore never drops a
*pinned* entry; of the un-pinned ones it keeps the ``keep`` most recently used per kind (train / eval), so alternating
between a full and a ragged batch shape does not re-allocate and re-zero gigabytes every epoch either.
"""
from __future__ import annotations

from collections import OrderedDict
from typing import Callable, Dict, Hashable, List

import torch


class WorkspaceCache:
    def __init__(self, keep: int = 2):
        self.keep = max(1, int(keep))
        self._entries: "OrderedDict[Hashable, torch.Tensor]" = OrderedDict()
        self._kind: Dict[Hashable, Hashable] = {}
        self._pins: Dict[Hashable, int] = {}

    def __len__(self) -> int:
        return len(self._entries)

    def __contains__(self, key) -> bool:
        return key in self._entries

    def keys(self):
        return list(self._entries.keys())

    def get(self, key: Hashable, kind: Hashable, alloc: Callable[[], torch.Tensor]) -> torch.Tensor:
        ws = self._entries.get(key)
        if ws is not None:
            self._entries.move_to_end(key)
            return ws
        # make room first (a workspace is GBs): oldest un-pinned entries of this kind beyond keep - 1
        same = [k for k in self._entries if self._kind[k] == kind and not self._pins.get(k)]
        for k in same[: max(0, len(same) - (self.keep - 1))]:
            del self._entries[k]
            del self._kind[k]
        ws = alloc()
        self._entries[key] = ws
        self._kind[key] = kind
        return ws

    def pin_all(self) -> List[Hashable]:
        """Pins every entry present now (call after the warm-up steps of a graph capture); returns the token for unpin."""
        token = list(self._entries.keys())
        for k in token:
            self._pins[k] = self._pins.get(k, 0) + 1
        return token

    def unpin(self, token) -> None:
        for k in token or ():
            c = self._pins.get(k, 0) - 1
            if c > 0:
                self._pins[k] = c
            else:
                self._pins.pop(k, None)

    def pinned(self, key) -> bool:
        return bool(self._pins.get(key))

    def clear(self) -> None:
        """Drops everything, pinned or not (the owner moved to another device: every recorded graph is invalid anyway)."""
        self._entries.clear()
        self._kind.clear()
        self._pins.clear()
