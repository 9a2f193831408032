"""NOVEL_PE renderer variant.

The reference's `src/models/novel_pe/pe_nerf_novel_renderer.py` is its NOVEL renderer with ONE line changed (:444: the
undeformed sample points are handed to the field as well); what differs is the field, `pe_novel_pixelnerf.py`, which
`diner_b200.novel.NeRFRendererDGS` recognises by the model's `deformation_layer` (see `diner_b200.novel`, `field.PackedPE`,
`dnr_encode_novel_pe_fwd`).  This module exists so that the NOVEL_PE YAMLs' `renderer.module` has a drop-in target:

    renderer.module: diner_b200.novel_pe.NeRFRendererDGS
"""
from .novel import NeRFRendererDGS  # noqa: F401
