"""B200-native MCMC log-posterior for PCA + Gaussian-process emulators (drop-in for jdmulligan/bayesian-inference's hot path)."""
__version__ = "0.1.0"
