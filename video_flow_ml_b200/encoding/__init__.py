from .flow_encoders import (FlowEncoder, FlowEncoderFactory, GamedevFlowEncoder, MotionVectorsRG8FlowEncoder,
                            decode_motion_vectors, encode_flow, encode_motion_vectors)

__all__ = ["FlowEncoder", "FlowEncoderFactory", "GamedevFlowEncoder", "MotionVectorsRG8FlowEncoder",
           "encode_flow", "encode_motion_vectors", "decode_motion_vectors"]
