"""Decoder factory with the interface of the reference's ``src/conv_onet/config.py:4-33``:
``get_model(cfg, nice)`` builds the NICE four-level decoder or the single iMAP* MLP from the
same yaml keys, so ``NICE_SLAM.__init__`` can call it unchanged."""
from .models import decoder_dict

# iMAP* hyper-parameters are constants of the reference factory, not yaml keys (config.py:29-32)
_IMAP_KW = dict(c_dim=0, color=True, hidden_size=256, skips=[], n_blocks=4)


def get_model(cfg, nice=True):
    embed = cfg['model']['pos_embedding_method']
    if not nice:
        return decoder_dict['imap'](dim=cfg['data']['dim'], pos_embedding_method=embed, **_IMAP_KW)
    lens = {f'{level}_grid_len': cfg['grid_len'][level] for level in ('coarse', 'middle', 'fine', 'color')}
    return decoder_dict['nice'](dim=cfg['data']['dim'], c_dim=cfg['model']['c_dim'], coarse=cfg['coarse'],
                                pos_embedding_method=embed, **lens)
