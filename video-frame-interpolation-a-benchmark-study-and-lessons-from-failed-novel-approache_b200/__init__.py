"""b200-vfi: B200-native RIFE v4.25 inference, AdaptiveVFI routing statistics and PSNR/SSIM, behind the
reference's models/base.py interface.  Import as `vfi_b200` (see /vfi_b200.py)."""
__version__ = "0.1.0"
PACKAGE_DIR = __path__[0]
