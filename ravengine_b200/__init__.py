"""ravengine_b200 — B200-native (sm_100a CUDA) render-prep path of RavEngine behind a C-ABI.

Stages: Hi-Z depth pyramid, instance cull + LOD + indirect compaction, compute skinning,
skinned draw-call prep, cluster grid build and light assignment. See DESIGN.md.
There is no CPU fallback: importing works anywhere, computing needs librve_b200.so + a GPU.
"""
from ._abi import (CullingUBO, CullingUBOinstance, GridAssignUBO, GridBuildUBO, IndirectIndexedCommand, SceneGenParams,
                   SkinningPrepareUBO, SkinningUBO, load_library)
from .api import Buffer, Context, Pyramid, RveError

__all__ = ["Context", "Buffer", "Pyramid", "RveError", "CullingUBO", "CullingUBOinstance", "GridAssignUBO", "GridBuildUBO",
           "IndirectIndexedCommand", "SceneGenParams", "SkinningPrepareUBO", "SkinningUBO", "load_library"]
