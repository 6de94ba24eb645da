"""Mirror of the reference ``yoho/src/config.py:4-29`` (``YOHOConfig``): the object
``get_batched_spectogram(config)`` receives.  Same field names and derived properties."""
from pydantic import BaseModel


class YOHOConfig(BaseModel):
    sample_rate: int
    n_mel_bands: int
    n_fft: int
    stft_hop: int

    max_audio_len: int
    max_text_len: int

    dims: int

    n_audio_heads: int
    n_audio_blocks: int

    n_text_heads: int
    n_text_blocks: int

    @property
    def n_samples(self) -> int:
        """Samples that give exactly ``max_audio_len`` frames (config.py:27-29)."""
        return self.n_fft + (self.max_audio_len - 1) * self.stft_hop

    @property
    def max_input_seconds(self) -> float:
        return self.n_samples / self.sample_rate
