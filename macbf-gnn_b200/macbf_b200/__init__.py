"""Host side of the B200 safety-certificate path: ctypes binding + torch tensor plumbing."""
from . import lib  # noqa: F401
