from typing import *  # noqa: F401,F403
