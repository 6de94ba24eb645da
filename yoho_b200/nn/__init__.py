from .whisper import (AudioEncoder, Whisper, conv_out_len, flatten_encoder_params, init_encoder_params,
                      sin_positional_encoding)
from . import model

__all__ = ["AudioEncoder", "Whisper", "conv_out_len", "flatten_encoder_params", "init_encoder_params",
           "sin_positional_encoding", "model"]
