"""Script 06's names, literally: `from vaeclust.nn06 import MLPVAE, vae_loss` replaces the model / loss definitions of
scripts/06_train_basic_vae_easy.py:145-188 (constructor order `(input_dim, hidden_dim, latent_dim)`, attributes
`enc1, enc2, mu, logvar, dec1, dec2, out`).  Script 19's differently-ordered `MLPVAE` lives in `vaeclust.nn`."""
from .nn import EasyMLPVAE as MLPVAE
from .nn import vae_loss

__all__ = ["MLPVAE", "vae_loss"]
