"""SageMaker configuration questionnaire (reference commands/config/sagemaker.py) — not available on this build:
the launcher drives local B200 nodes (torchrun / simple launcher); there is no managed-cloud back-end."""


def get_sagemaker_input():
    raise NotImplementedError("Amazon SageMaker is not a supported compute environment of the B200-native build; "
                              "choose `This machine` in `accelerate config`.")
