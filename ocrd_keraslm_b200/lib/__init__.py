"""backend library interface (mirrors ocrd_keraslm.lib, /root/reference/ocrd_keraslm/lib/__init__.py:7)

Rater - LM application on the B200 rating path
Node - tree data type for beam search
"""
from .rating import Rater, Node, DeviceState
