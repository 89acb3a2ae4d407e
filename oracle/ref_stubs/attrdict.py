"""Stub so that the read-only reference (`mantrap/environment/sgan.py:5`) imports in this container.
Test infrastructure only (oracle/): never imported by the product package."""


class AttrDict(dict):
    __getattr__ = dict.__getitem__
    __setattr__ = dict.__setitem__
