"""B200-native (sm_100a) hot path of ishine/pix2pixHDAudioSR: MDCT4/IMDCT4 and the pix2pixHD conv
GAN step behind the reference's own Python API.  Device code lives in csrc/ behind the C ABI of
include/p2p_b200.h; this package is the host-side mirror of the reference interface."""
__version__ = "0.1.0"
