"""The X-engine blocks, importable under the reference's names (lwa352_pipeline.blocks.*)."""
from .block_base import Block  # noqa: F401
from .corr_block import Corr  # noqa: F401
from .corr_acc_block import CorrAcc  # noqa: F401
from .corr_subsel_block import CorrSubsel  # noqa: F401
from .corr_output_full_block import CorrOutputFull  # noqa: F401
from .corr_output_part_block import CorrOutputPart  # noqa: F401
from .beamform_block import Beamform  # noqa: F401
from .beamform_sum_beams_block import BeamformSumBeams  # noqa: F401
from .beamform_sum_block import BeamformSum  # noqa: F401
from .beamform_sum_single_beam_block import BeamformSumSingleBeam  # noqa: F401
from .beamform_output_block import BeamformOutput  # noqa: F401
from .beamform_vlbi_output_block import BeamformVlbiOutput  # noqa: F401
from .copy_block import Copy  # noqa: F401
from .dummy_source_block import DummySource  # noqa: F401
