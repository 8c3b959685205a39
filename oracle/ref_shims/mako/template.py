"""Import shim for mako.template: constructed at class-definition time, never rendered here."""


class Template:
    def __init__(self, text="", **kwargs):
        self.source = text

    def render(self, **kwargs):
        raise RuntimeError("mako shim: rendering is not available (gt:* backends cannot build here)")
