"""Error convention of the reference (ipex_llm/utils/common/log4Error.py:22-32):
``invalidInputError(cond, msg, fix)`` logs a banner and raises ``RuntimeError(msg)``."""
import logging

logger = logging.getLogger("ipex_llm_b200")


def invalidInputError(condition, errMsg, fixMsg=None):
    if not condition:
        logger.error("\n\n****************************Usage Error************************\n" + errMsg)
        if fixMsg:
            logger.error("\n\n**************************How to fix***********************\n" + fixMsg)
        raise RuntimeError(errMsg)


def invalidOperationError(condition, errMsg, fixMsg=None, cause=None):
    if not condition:
        logger.error("\n\n****************************Usage Error************************\n" + errMsg)
        if cause:
            raise cause
        raise RuntimeError(errMsg)
