"""B200-native (sm_100a) implementation of the DefectGAN generator / discriminator training hot path of
jason2714/de-i2i-gan.  Python here is host plumbing; all device work goes through libdei2i.so (include/dei2i.h)."""
__version__ = "0.1.0"
