"""Exit-code classification for restartPolicy: ExitCode
(reference pkg/util/train/train_util.go:18-53)."""

PERMANENT_EXIT_CODES = frozenset({1, 2, 126, 127, 128, 139})
RETRYABLE_SIGNAL_EXIT_CODES = frozenset({130, 137, 143})
USER_RETRYABLE_EXIT_CODE = 138  # 128 + SIGUSR1: "please restart me"


def is_retryable_exit_code(exit_code: int) -> bool:
    if exit_code in PERMANENT_EXIT_CODES:
        return False
    if exit_code in RETRYABLE_SIGNAL_EXIT_CODES:
        return True
    if exit_code == USER_RETRYABLE_EXIT_CODE:
        return True
    return False  # anything else is treated as permanent
