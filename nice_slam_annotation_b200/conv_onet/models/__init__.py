"""Mirror of src/conv_onet/models/__init__.py:4-7."""
from . import decoder

decoder_dict = {
    'nice': decoder.NICE,
    'imap': decoder.MLP,
}
