"""HuggingFace integrations (reference smdistributed/modelparallel/torch/nn/huggingface/)."""
