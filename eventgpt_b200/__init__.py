"""eventgpt_b200: the B200 (sm_100a) event -> visual-token front end of EventGPT.

``common``  mirrors the reference's ``common/common.py`` (event loading, frame
            splitting, rasterisation, CLIP preprocessing);
``model``   mirrors the visual half of ``model/EventChatModel.py`` (tower,
            projector, adaptor, spatio-temporal pooling);
``events``  the device-resident operators underneath;
``_lib``    ctypes binding of libegpt_b200.so (include/egpt_b200.h).
There is no CPU fallback anywhere in this package.
"""
from . import _lib  # noqa: F401

__all__ = ["_lib", "common", "events", "model", "weights", "sharding"]
