/* stub: nothing from this header is used by old/chaos_game.c */
