"""Launched end-to-end scripts that exercise whole features (metrics gathering, checkpoint resume, convergence, memory, pipeline
inference, DeepSpeed-style configurations).  In the reference these need external packages (datasets, evaluate, transformers,
deepspeed); here they run on the native model families with synthetic data, so they work offline and on CPU."""
