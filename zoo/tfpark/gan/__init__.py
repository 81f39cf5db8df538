from .common import (minimax_discriminator_loss, minimax_generator_loss, modified_discriminator_loss, modified_generator_loss,  # noqa: F401
                     wasserstein_discriminator_loss, wasserstein_generator_loss)
from .gan_estimator import GANEstimator  # noqa: F401
