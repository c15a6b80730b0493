"""Model zoo: benchmark models and Trial definitions built on the sm_100a ops."""
