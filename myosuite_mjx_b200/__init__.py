"""B200-native batched musculoskeletal step behind the MJX-style API of micropilot/myosuite-mjx."""
from . import mjcf  # noqa: F401
