"""dvs_of_training_framework_b200 -- the B200 (sm_100a) hot path of
e-sha/dvs_of_training_framework behind the reference's own Python signatures.

    encoder.compute_event_image / quantize / extend_size   <- EV_FlowNet/net.py:88-205
    loss.Loss / Losses / init_losses / charbonier_loss      <- utils/loss.py
    data.get_count_image / EventCrop / RandomRotation /
         frame_generator                                    <- utils/data.py
    transformation.map                                      <- utils/transformation (native ext)
    eval.flow_error_dense                                   <- utils/eval.py:6-50
    dataset.encode_batch / decode_batch (+ _info)          <- utils/dataset.py:201-373
    dataset.collate_wrapper                                 <- utils/dataset.py:961-1020
    preprocessed.PreprocessedDataloader / select_*_ranges /
         join_batches / read_encoded_batch                  <- utils/dataset.py:19-198,400-425,505-549,799-954

Everything computes in libdvsof.so (hand-written CUDA, C ABI in include/dvsof.h); there is no
CPU fallback."""
from . import _lib  # noqa: F401
from .encoder import (compute_event_image, compute_event_image_compact, compute_voxel_grid, extend_size,  # noqa: F401
                      quantize)
from .loss import Loss, Losses, charbonier_loss, init_losses, interpolate  # noqa: F401
from .data import (EventCrop, RandomRotation, augment_events, frame_generator, frame_offsets,  # noqa: F401
                   get_count_image, rotation_lut)
from . import transformation  # noqa: F401
from . import preprocessed  # noqa: F401
from .preprocessed import PreprocessedDataloader, join_batches, select_encoded_ranges  # noqa: F401
from .eval import flow_error_dense, flow_error_dense_device  # noqa: F401
from .dataset import (collate_wrapper, decode_batch, decode_batch_info, decode_events, decode_quantized_batch,  # noqa: F401
                      encode_batch, encode_batch_info, encode_quantized_batch, quantize_batch)

__all__ = ['compute_event_image', 'compute_event_image_compact', 'compute_voxel_grid', 'extend_size', 'quantize',
           'Loss', 'Losses', 'charbonier_loss', 'init_losses', 'interpolate',
           'EventCrop', 'RandomRotation', 'get_count_image', 'frame_generator', 'frame_offsets', 'augment_events', 'rotation_lut',
           'transformation', 'flow_error_dense', 'flow_error_dense_device',
           'collate_wrapper', 'encode_batch', 'decode_batch', 'encode_batch_info', 'decode_batch_info', 'decode_events',
           'quantize_batch', 'encode_quantized_batch', 'decode_quantized_batch']
