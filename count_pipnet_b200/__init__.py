"""count_pipnet_b200: B200 (sm_100a) forward hot path of PIP-Net / Count-PIP-Net behind the reference's nn.Module API."""
__version__ = "0.1.0"
