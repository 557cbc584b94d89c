/* TEST INFRASTRUCTURE ONLY (oracle/): demod_binary.c:46 includes a generated header that is not in
 * the reference tree (debian/patches/define_git_version.patch creates the same one-liner). */
#define ERP_GIT_VERSION "0"
