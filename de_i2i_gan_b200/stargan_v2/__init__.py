"""Host-side mirror of the reference's ``stargan-v2`` fork for the hot path (adain mode, w_hpf = 0): ``core.model`` network
classes and the ``core.solver`` loss functions on the sm_100a kernels."""
