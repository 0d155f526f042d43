"""tilelang/carver/arch/driver/cuda_driver.py: the device queries GEMM examples make."""
import torch


def get_num_sms(device: int = 0) -> int:
    if not torch.cuda.is_available():
        raise RuntimeError("get_num_sms: no CUDA device (the B200 backend has no CPU path)")
    return torch.cuda.get_device_properties(device).multi_processor_count


def get_cuda_device_properties(device: int = 0):
    return torch.cuda.get_device_properties(device)
