"""`python main_part3.py ...` -- same command line as the reference's main_part3.py; the work happens in
distributed_pytorch_b200.cli.run("part3")."""
from distributed_pytorch_b200.cli import main

if __name__ == "__main__":
    main("part3")
