/* TEST INFRASTRUCTURE ONLY (oracle/): command-line entry for the reference driver built by
 * oracle/Makefile.  Equivalent to the reference's own standalone wrapper (demod_binary.c:1786-1794),
 * which is renamed at compile time so the same objects can also be put in a shared library. */
#ifdef __cplusplus
extern "C"
#endif
int MAIN(int argc, char *argv[]);
int main(int argc, char *argv[]) { return MAIN(argc, argv); }
