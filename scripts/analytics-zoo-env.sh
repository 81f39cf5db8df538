# Reference name of zoo-env.sh.
. "$(cd "$(dirname "${BASH_SOURCE[0]}")" && pwd)/zoo-env.sh"
