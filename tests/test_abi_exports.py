"""CPU checks of the drop-in boundary: the C-ABI library loads without a GPU, exports every symbol that
include/*.h declares, the ctypes mirror has a prototype for each, the POD layouts have the reference's
sizes/offsets, and compute entry points fail loudly (no CPU fallback) when there is no device."""
import ctypes as C
import re
from pathlib import Path

import pytest

from ravengine_b200 import _abi

ROOT = Path(__file__).resolve().parent.parent
DECL = re.compile(r"RVE_API\s+[\w\s\*]+?\b(r(?:ve|vs)_\w+)\s*\(")


def declared_symbols():
    names = []
    for h in ("rve_b200.h", "rve_shim.h"):
        names += DECL.findall((ROOT / "include" / h).read_text())
    return sorted(set(names))


def test_headers_declare_something():
    names = declared_symbols()
    assert len(names) >= 50
    for must in ("rve_cull", "rve_depth_pyramid_build", "rve_skin", "rve_skinned_drawcall_prepare", "rve_cluster_build",
                 "rve_cluster_assign", "rvs_cull_static", "rvs_cull_skeletal", "rvs_bin_lights"):
        assert must in names


def test_library_exports_every_declared_symbol():
    lib = _abi.load_library()
    for name in declared_symbols():
        assert hasattr(lib, name), f"{name} declared in include/*.h but not exported by librve_b200.so"


def test_ctypes_mirror_covers_every_declared_symbol():
    missing = [n for n in declared_symbols() if n not in _abi.PROTOTYPES]
    assert not missing, f"no ctypes prototype for {missing}"
    extra = [n for n in _abi.PROTOTYPES if n not in declared_symbols()]
    assert not extra, f"ctypes prototypes without a header declaration: {extra}"


def test_every_declaration_cites_the_reference():
    """Each entry point documents the reference interface it replaces (file:line)."""
    text = (ROOT / "include" / "rve_b200.h").read_text()
    for stage, cite in [("rve_cull", "defaultcull.csh"), ("rve_skin", "skinning_cs.csh"),
                        ("rve_skinned_drawcall_prepare", "skinned_mesh_drawcall.csh"), ("rve_cluster_build", "cluster_build_grid.csh"),
                        ("rve_cluster_assign", "cluster_assign_lights.csh"), ("rve_depth_pyramid_build", "depthpyramid")]:
        pos = text.index(f" {stage}(")
        assert cite in text[max(0, pos - 1500):pos], f"{stage} does not cite {cite}"


@pytest.mark.parametrize("struct,size,offsets", [
    (_abi.IndirectIndexedCommand, 20, {"indexCount": 0, "instanceCount": 4, "indexStart": 8, "baseVertex": 12, "baseInstance": 16}),
    (_abi.CullingUBO, 88, {"viewProj": 0, "camPos": 64, "numCubos": 76, "cameraRenderLayers": 80, "singleInstanceModeAndShadowMode": 84}),
    (_abi.CullingUBOinstance, 36, {"indirectBufferOffset": 0, "numObjects": 4, "cullingBufferOffset": 8, "radius": 12, "numLODs": 16,
                                   "idOutputBufferBindlessHandle": 20, "entityIDInputBufferBindlessHandle": 24,
                                   "indirectOutputBufferBindlessHandle": 28, "lodDistanceBufferBindlessHandle": 32}),
    (_abi.SkinningUBO, 24, {"numObjects": 0, "numVertices": 4, "numBones": 8, "boneReadOffset": 12, "vertexWriteOffset": 16, "vertexReadOffset": 20}),
    (_abi.SkinningPrepareUBO, 28, {"indexBufferOffset": 0, "vertexBufferOffset": 4, "nIndicesInThisMesh": 8, "nVerticesInThisMesh": 12,
                                   "nTotalObjects": 16, "drawCallBufferOffset": 20, "baseInstanceOffset": 24}),
    (_abi.GridBuildUBO, 92, {"invProj": 0, "gridSize": 64, "zNear": 76, "screenDim": 80, "zFar": 88}),
    (_abi.GridAssignUBO, 72, {"viewMat": 0, "pointLightCount": 64, "spotLightCount": 68}),
])
def test_pod_layouts_match_reference(struct, size, offsets):
    # sizes and offsets of the reference's push-constant / storage structs
    assert C.sizeof(struct) == size
    for field, off in offsets.items():
        assert getattr(struct, field).offset == off, (struct.__name__, field)


def test_header_static_asserts_compile_as_c_and_cpp(tmp_path):
    """The header is plain C and its layout asserts hold for the host compiler too."""
    import subprocess

    src = tmp_path / "t.c"
    src.write_text('#include "rve_b200.h"\n#include "rve_shim.h"\nint main(void){return sizeof(RveCluster)==160 && sizeof(RvePointLight)==512 '
                   '&& sizeof(RveSpotLight)==168 && sizeof(RveVertexJointBinding)==32 ? 0 : 1;}\n')
    for cc, std in (("gcc", "-std=c11"), ("g++", "-std=c++17")):
        exe = tmp_path / f"t_{cc}"
        subprocess.run([cc, std, "-x", "c" if cc == "gcc" else "c++", f"-I{ROOT / 'include'}", str(src), "-o", str(exe)], check=True)
        assert subprocess.run([str(exe)]).returncode == 0


def test_no_device_fails_loudly():
    """Without a CUDA device the context cannot be created: there is no CPU fallback behind the ABI."""
    import torch

    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    lib = _abi.load_library()
    h = C.c_void_p()
    rc = lib.rve_ctx_create(0, None, C.byref(h))
    assert rc == -5  # RVE_ERR_NO_DEVICE
    assert b"no CPU fallback" in lib.rve_last_error_string()
    from ravengine_b200 import Context, RveError

    with pytest.raises(RveError):
        Context(0)


def test_product_never_imports_the_oracle():
    """oracle/ is test infrastructure: nothing under ravengine_b200/ may import, load or link it."""
    for f in (ROOT / "ravengine_b200").rglob("*"):
        if f.suffix in (".py", ".cu", ".h", ".cpp"):
            t = f.read_text()
            assert "import oracle" not in t and "from oracle" not in t and "liboracle" not in t and "oracle.h" not in t, f
