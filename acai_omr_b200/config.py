"""Special-token constants of the LMX vocabulary (reference acai_omr/config.py:12-14)."""
LMX_BOS_TOKEN = "<bos>"
LMX_EOS_TOKEN = "<eos>"
LMX_PAD_TOKEN = "<pad>"  # token used for padding lmx sequences
