class WebsocketSession:
    def __init__(self, *a, **kw):
        pass
