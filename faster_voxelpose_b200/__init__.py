"""faster_voxelpose_b200 -- B200-native (sm_100a) voxel projection hot path of Faster-VoxelPose.

``lib/`` mirrors the reference's ``lib/`` module names (``models.project_whole``,
``models.project_individual``, ``core.proposal``, ``models.human_detection_net``,
``models.joint_localization_net``, ``models.faster_voxelpose``); ``csrc/`` holds the CUDA kernels
behind the C ABI of ``include/fvp_b200.h``; ``ops`` is the tensor-level binding.
"""
__version__ = "0.1.0"
