from .whisper import (AudioEncoder, TextDecoder, Whisper, conv_out_len, cross_attention_kv, flatten_cross_kv_params,
                      flatten_decoder_params, flatten_encoder_params, init_decoder_params, init_encoder_params,
                      sin_positional_encoding)
from . import model

__all__ = ["AudioEncoder", "TextDecoder", "Whisper", "conv_out_len", "cross_attention_kv", "flatten_cross_kv_params",
           "flatten_decoder_params", "flatten_encoder_params", "init_decoder_params", "init_encoder_params",
           "sin_positional_encoding", "model"]
