"""lomond stand-in: the WebSocket client is never used by the local benchmark path."""


class WebSocket:
    def __init__(self, *a, **kw):
        raise RuntimeError("lomond is not installed: master connections are unavailable in the reference arm")


class events:  # noqa: N801
    class Event: pass
    class Binary(Event): pass
    class Text(Event): pass
    class Ready(Event): pass
    class Poll(Event): pass
    class Closed(Event): pass
    class Disconnected(Event): pass
    class ConnectFail(Event): pass
    class Rejected(Event): pass
    class BackOff(Event): pass
    class Connecting(Event): pass
    class Connected(Event): pass
    class Closing(Event): pass
    class Ping(Event): pass
    class Pong(Event): pass
    class ProtocolError(Event): pass
    class UnknownMessage(Event): pass
    class Unresponsive(Event): pass
