from .flow_encoders import (FlowEncoder, FlowEncoderFactory, GamedevFlowEncoder, HSVFlowEncoder, MotionVectorsRG8FlowEncoder,
                            MotionVectorsRGB8FlowEncoder, TorchvisionFlowEncoder, decode_motion_vectors, encode_flow,
                            encode_motion_vectors)

__all__ = ["FlowEncoder", "FlowEncoderFactory", "GamedevFlowEncoder", "HSVFlowEncoder", "MotionVectorsRG8FlowEncoder",
           "MotionVectorsRGB8FlowEncoder", "TorchvisionFlowEncoder", "encode_flow", "encode_motion_vectors",
           "decode_motion_vectors"]
