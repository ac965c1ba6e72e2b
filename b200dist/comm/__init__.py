"""Peer-memory communication layer (NVLink 5 / NVSwitch): symmetric buffers + device-side collectives."""
from .collectives import Communicator, get_communicator
from .peer_mem import PeerBuffer

__all__ = ["Communicator", "get_communicator", "PeerBuffer"]
