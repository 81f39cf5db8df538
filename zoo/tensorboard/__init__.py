from .writer import Summary, TrainSummary, ValidationSummary, InferenceSummary, FileWriter, FileReader  # noqa: F401
