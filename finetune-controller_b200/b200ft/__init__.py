"""B200-native fine-tuning step: host-side Python over the libb200ft C ABI (include/b200ft.h)."""
__all__ = ["ops", "gpt2"]
