"""lvc_b200 -- B200-native clip construction and patch embedding."""
__version__ = "0.1.0"
