/* Stub of the Sunway vendor header <crts.h>; intentionally empty (see athread.h). */
