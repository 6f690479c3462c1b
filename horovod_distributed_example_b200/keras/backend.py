"""The handful of ``keras.backend`` calls the reference makes (``mnist_keras.py:36,50,122``)."""
from .models import clear_session  # noqa: F401

_learning_phase = 1


from .layers import image_data_format, set_image_data_format  # noqa: F401,E402


def set_learning_phase(value: int) -> None:
    global _learning_phase
    _learning_phase = int(value)


def learning_phase() -> int:
    return _learning_phase


def set_session(session) -> None:      # TF1 session plumbing: nothing to do here
    return None


def epsilon() -> float:
    return 1e-7
