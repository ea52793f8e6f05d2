"""Import surface of the reference kept as is: `from shellac.server import HttpParser, StreamBuf`
(/root/reference/src/python/shellac/server/__init__.py:5-6) becomes
`from shellac_b200.server import HttpParser, StreamBuf`."""
from .HttpParser import HttpParser
from .StreamBuf import StreamBuf

__all__ = ['HttpParser', 'StreamBuf']
