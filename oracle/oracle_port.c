/* placeholder, replaced below */
int oracle_port_placeholder(void) { return 0; }
