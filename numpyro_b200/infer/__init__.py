from .hmc import HMC, NUTS, HMCAdaptState, HMCState, MCMCKernel, init_to_uniform, init_to_value
from .hmc_gibbs import HMCECS, HMCECSState, HMCGibbs, HMCGibbsState
from .mcmc import MCMC
from .util import Predictive, log_likelihood

__all__ = ["HMC", "NUTS", "MCMC", "MCMCKernel", "HMCState", "HMCAdaptState", "init_to_uniform",
           "init_to_value", "Predictive", "log_likelihood", "HMCECS", "HMCGibbs", "HMCECSState",
           "HMCGibbsState"]
