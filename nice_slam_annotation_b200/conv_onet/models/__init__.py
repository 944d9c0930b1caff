"""Decoder registry, keyed as in the reference (src/conv_onet/models/__init__.py:4-7)."""
from .decoder import MLP, NICE

decoder_dict = dict(nice=NICE, imap=MLP)
