// placeholder, replaced below
extern "C" int sgx_oracle_abi_version() { return 1; }
