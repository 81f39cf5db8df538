from .knrm import KNRM  # noqa: F401
from .text_matcher import TextMatcher  # noqa: F401
