"""``EasyDict``: dict with attribute access — the result type of the reference decoders
(main/dnnlib/util.py:42-54)."""
from typing import Any


class EasyDict(dict):
    def __getattr__(self, name: str) -> Any:
        try:
            return self[name]
        except KeyError:
            raise AttributeError(name)

    def __setattr__(self, name: str, value: Any) -> None:
        self[name] = value

    def __delattr__(self, name: str) -> None:
        del self[name]
