"""Inert stand-in for gradio (web_ui.py:17, 224-235): records what the reference builds, serves nothing."""


class Image:
    def __init__(self, *args, **kwargs):
        self.kwargs = kwargs


class Interface:
    def __init__(self, fn=None, inputs=None, outputs=None, **kwargs):
        self.fn, self.inputs, self.outputs, self.kwargs = fn, inputs, outputs, kwargs

    def launch(self, *args, **kwargs):
        raise RuntimeError("the gradio shim does not serve")
