#define GIT_VERSION "ref"
#define GIT_HASH "ref"
#define GIT_BRANCH "ref"
