from .audio import (get_batched_spectogram, log_mel_spectogram, log_mel_spectogram_chunked, mel_spectogram,
                    normalize_spectogram, whisper_normalize_spectogram)
from .mel_filterbanks import mel_filter_banks

__all__ = ["mel_spectogram", "get_batched_spectogram", "normalize_spectogram", "whisper_normalize_spectogram",
           "log_mel_spectogram", "log_mel_spectogram_chunked", "mel_filter_banks"]
