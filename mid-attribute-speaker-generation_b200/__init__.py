"""B200-native batched synthesis path (FastSpeech2 + TacoSpawn GMM prior + HiFi-GAN)."""
from . import synthetic  # noqa: F401
