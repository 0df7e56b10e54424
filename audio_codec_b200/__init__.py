"""B200-native batched LC3plus encoder (and error-free decoder).

The product is `liblc3plus_b200.so` (C ABI in include/lc3plus_b200.h: the ETSI reference encoder API plus
`lc3plus_enc_batch`), built from audio_codec_b200/csrc.  This package is the thin Python mirror of that C
API (ctypes) used by the tests and bench.py; there is no Python or CPU encode path.
"""
from .api import (LC3Error, Encoder, EncoderArray, enc_batch, kernel_launches, last_error, load_library, library_path,  # noqa: F401
                  profile_enable, profile_read, set_chunk_frames, container_header, container_frames,
                  write_bitstream_file, read_switching_file, enc_fresh_noalloc, set_device, sync, profile_read_device, set_bitrate_many,
                  Decoder, DecoderArray, dec_batch, dec_kernel_launches, dec_profile_enable, dec_profile_read)
